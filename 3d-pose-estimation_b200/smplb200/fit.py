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
        self.loss = z(3)
        L = _lib.lib()
        self.scratch = torch.empty(L.smplb_fit_scratch_bytes(B), dtype=torch.uint8, device=dev)
        cfg = dict(DEFAULT_CAM)
        cfg.update(cam_cfg or {})
        scale = cfg['ss'] if reproj_scale is None else reproj_scale          # config.py:22: 1.0 * ss
        self.losscfg = LossCfg(CamCfg(**cfg), scale, reproj_cap or 0.0, 2.0, 1.0)
        self.n_total = int(n_total if n_total is not None else B * NJ * 2)
        self.hyper = (lr, beta1, beta2, eps)
        self.t = 0

    def _io(self, step):
        p = lambda t: None if t is None else C.c_void_p(t.data_ptr())
        lr, b1, b2, eps = self.hyper
        return FitIO(self.B, p(self.beta), p(self.theta), p(self.cam_pos), p(self.cam_rot), p(self.focal),
                     p(self.gt_px), p(self.px), p(self.g_px), p(self.g_beta), p(self.g_theta), p(self.m_theta),
                     p(self.v_theta), p(self.m_beta), p(self.v_beta), p(self.loss), p(self.scratch), self.losscfg,
                     self.n_total, step, lr, b1, b2, eps)

    def step(self):
        """One fused fitting step (5-6 kernel launches on the current stream).  Returns the device loss tensor
        [uncapped share, capped, gate] without synchronising."""
        self.t += 1
        io = self._io(self.t)
        with torch.cuda.device(self.dev):
            check(_lib.lib().smplb_fit_step(self.model._h, C.byref(io),
                                            C.c_void_p(torch.cuda.current_stream(self.dev).cuda_stream)))
        return self.loss
