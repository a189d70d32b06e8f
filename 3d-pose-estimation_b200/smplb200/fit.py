"""Iterative SMPL fitting on the 2D re-projection loss (BASELINE config 3): host-side driver of smplb_fit_step.

Per-sample pose (optionally shape) are the Adam variables (tf.train.AdamOptimizer formula,
pose_model_3d.py:393-401); the loss is the re-projection block of pose_model_3d.py:194-238.  All arithmetic is
in libsmplb200.so; torch tensors are carriers.  Multi-GPU: every rank fits its own batch shard, the only
collective is the all-reduce of the loss scalars (see dist.py)."""
import ctypes as C

import torch

from . import _lib
from ._lib import NJ, NB, FitIO, LossCfg, CamCfg, check
from .api import DEFAULT_CAM, _dev_f32


class Fitter:
    def __init__(self, model, beta, theta, cam_pos, cam_rot, gt_px, focal=None, fit_beta=False, lr=2e-4,
                 beta1=0.9, beta2=0.999, eps=1e-8, reproj_scale=None, reproj_cap=0.0, n_total=None, cam_cfg=None):
        """beta [B,10], theta [B,72] (initial values, copied), camera [B,3],[B,3],([B]), gt_px [B,24,2] pixels.
        reproj_cap=1/24 reproduces the reference's tf.minimum cap (:237); 0 disables it.
        n_total: element count of the global mean (global_B*24*2) when the batch is sharded over ranks."""
        dev = model.device
        self.model, self.dev = model, dev
        self.beta = _dev_f32(beta, dev).clone()
        self.theta = _dev_f32(theta, dev).clone()
        B = self.B = self.beta.shape[0]
        self.cam_pos = _dev_f32(cam_pos, dev, (B, 3))
        self.cam_rot = _dev_f32(cam_rot, dev, (B, 3))
        self.focal = None if focal is None else _dev_f32(focal, dev, (B,))
        self.gt_px = _dev_f32(gt_px, dev, (B, NJ, 2))
        z = lambda *s: torch.zeros(s, dtype=torch.float32, device=dev)
        self.px, self.g_px = z(B, NJ, 2), z(B, NJ, 2)
        self.g_beta, self.g_theta = z(B, NB), z(B, NJ * 3)
        self.m_theta, self.v_theta = z(B, NJ * 3), z(B, NJ * 3)
        self.m_beta, self.v_beta = (z(B, NB), z(B, NB)) if fit_beta else (None, None)
        self.loss = z(4)                                   # [share of the mean, capped, gate, Adam lr_t]
        # Adam's step counter lives on the device (smplb_fit_io.step_dev): a captured step advances it on every replay
        self.step_dev = torch.zeros(1, dtype=torch.int32, device=dev)
        L = _lib.lib()
        self.scratch = torch.empty(L.smplb_fit_scratch_bytes(B), dtype=torch.uint8, device=dev)
        cfg = dict(DEFAULT_CAM)
        cfg.update(cam_cfg or {})
        scale = cfg['ss'] if reproj_scale is None else reproj_scale          # config.py:22: 1.0 * ss
        self.losscfg = LossCfg(CamCfg(**cfg), scale, reproj_cap or 0.0, 2.0, 1.0)
        self.n_total = int(n_total if n_total is not None else B * NJ * 2)
        self.hyper = (lr, beta1, beta2, eps)
        self.t = 0

    def _io(self, step, loss):
        p = lambda t: None if t is None else C.c_void_p(t.data_ptr())
        lr, b1, b2, eps = self.hyper
        return FitIO(self.B, p(self.beta), p(self.theta), p(self.cam_pos), p(self.cam_rot), p(self.focal),
                     p(self.gt_px), p(self.px), p(self.g_px), p(self.g_beta), p(self.g_theta), p(self.m_theta),
                     p(self.v_theta), p(self.m_beta), p(self.v_beta), p(loss), p(self.scratch), self.losscfg,
                     self.n_total, step, lr, b1, b2, eps, p(self.step_dev))

    def step(self, loss_out=None):
        """One fused fitting step (5-6 kernel launches on the current stream).  Returns the device loss tensor
        [uncapped share, capped, gate, lr_t] without synchronising.  loss_out: another 4-float device tensor to
        receive them (e.g. a row of a per-step log that is all-reduced once every k steps)."""
        self.t += 1                    # host mirror only; the kernels read the device counter
        loss = self.loss if loss_out is None else loss_out
        if loss.numel() < 4 or loss.dtype != torch.float32 or not loss.is_contiguous():
            raise ValueError('loss_out must be a contiguous float32 tensor of >= 4 elements')
        io = self._io(self.t, loss)
        with torch.cuda.device(self.dev):
            check(_lib.lib().smplb_fit_step(self.model._h, C.byref(io),
                                            C.c_void_p(torch.cuda.current_stream(self.dev).cuda_stream)))
        return loss

    def reset_optimizer(self):
        """Zero the Adam moments and the device step counter (stream-ordered)."""
        self.t = 0
        self.step_dev.zero_()
        for t in (self.m_theta, self.v_theta, self.m_beta, self.v_beta):
            if t is not None:
                t.zero_()


class MeshFitter:
    """Iterative fitting on the reference's 3D losses (pose_model_3d.py:183-192: mesh_loss = MSE(verts)*1.0 +
    joint_loss = MSE(J_transformed)*2.0) with per-sample pose (optionally shape) as the Adam variables -- the variant
    of BASELINE config 3 that drives the backward through the VERTICES (skinning adjoint, g_A reduction, transposed
    blend GEMM).  One step = smplb_forward(KEEP_VPOSED) -> smplb_mse_grad x2 -> smplb_backward -> smplb_adam_step,
    all on the current stream; buffers are allocated once."""

    def __init__(self, model, beta, theta, gt_verts, gt_joints, fit_beta=False, lr=2e-4, beta1=0.9, beta2=0.999,
                 eps=1e-8, mesh_scale=1.0, joint_scale=2.0, n_total_verts=None, n_total_joints=None):
        from .api import _dev_f32 as f32
        dev = model.device
        self.model, self.dev = model, dev
        self.beta, self.theta = f32(beta, dev).clone(), f32(theta, dev).clone()
        B = self.B = self.beta.shape[0]
        self.gt_verts, self.gt_joints = f32(gt_verts, dev, (B, 6890, 3)), f32(gt_joints, dev, (B, NJ, 3))
        z = lambda *s: torch.zeros(s, dtype=torch.float32, device=dev)
        self.g_verts, self.g_joints = z(B, 6890, 3), z(B, NJ, 3)
        self.m_theta, self.v_theta = z(B, NJ * 3), z(B, NJ * 3)
        self.m_beta, self.v_beta = (z(B, NB), z(B, NB)) if fit_beta else (None, None)
        self.loss = z(1)
        self.step_dev = torch.zeros(1, dtype=torch.int32, device=dev)    # device-side Adam step counter (graph-safe)
        self.lr_t = z(1)
        self.out = {}
        self.scratch = torch.empty(int(_lib.lib().smplb_mse_scratch_bytes()), dtype=torch.uint8, device=dev)
        self.scales = (mesh_scale, joint_scale)
        self.n_total = (int(n_total_verts or B * 6890 * 3), int(n_total_joints or B * NJ * 3))
        self.hyper = (lr, beta1, beta2, eps)
        self.t = 0

    def step(self):
        """Returns the device loss tensor [1] (this shard's share of mesh_loss + joint_loss) without synchronising."""
        from . import api
        self.t += 1
        m = self.model
        o = m.forward(self.beta, self.theta, want_kp=False, want_Rs=False, flags=_lib.F_KEEP_VPOSED, out=self.out)
        api.mse_grad(o['verts'], self.gt_verts, self.scales[0], self.n_total[0], loss_out=self.loss, g_out=self.g_verts,
                     scratch=self.scratch)
        api.mse_grad(o['joints'], self.gt_joints, self.scales[1], self.n_total[1], loss_out=self.loss, accumulate=True,
                     g_out=self.g_joints, scratch=self.scratch)
        g_beta, g_theta = m.backward(self.beta, self.theta, g_verts=self.g_verts, g_joints=self.g_joints)
        lr, b1, b2, eps = self.hyper
        api.adam_tick(self.step_dev, self.lr_t, lr, b1, b2)
        api.adam_step_dev(self.theta, g_theta, self.m_theta, self.v_theta, self.lr_t, b1, b2, eps)
        if self.m_beta is not None:
            api.adam_step_dev(self.beta, g_beta, self.m_beta, self.v_beta, self.lr_t, b1, b2, eps)
        return self.loss
