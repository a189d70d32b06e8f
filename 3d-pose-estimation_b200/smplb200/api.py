"""Host-side wrapper over the C ABI.  torch CUDA tensors are only carriers (device memory +
streams); all arithmetic happens in libsmplb200.so.
"""
import ctypes as C
import pickle

import numpy as np
import torch

from . import _lib
from ._lib import NV, NV3, NJ, NB, ForwardIO, BackwardIO, CamCfg, LossCfg, check

DEFAULT_CAM = dict(img_w=320.0, img_h=240.0, ss=0.024, focal_const=0.05)   # config.py:9-12


def _ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _stream(device):
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def _dev_f32(x, device, shape=None):
    """Carrier: contiguous fp32 CUDA tensor on `device` (accepts numpy / torch / lists)."""
    if not torch.is_tensor(x):
        x = torch.as_tensor(np.asarray(x, dtype=np.float32))
    x = x.to(device=device, dtype=torch.float32).contiguous()
    if shape is not None and tuple(x.shape) != tuple(shape):
        raise ValueError('expected shape %s, got %s' % (tuple(shape), tuple(x.shape)))
    return x


def load_pkl(pkl_path):
    """Same loader contract as SMPL.__init__ (batch_smpl.py:22-23)."""
    with open(pkl_path, 'rb') as f:
        return pickle.load(f, encoding='latin1')


class Model:
    """Device-resident SMPL constants (SMPL.__init__, deps/tf_smpl/batch_smpl.py:17-69)."""

    def __init__(self, pkl, joint_type='cocoplus', device=None):
        dd = pkl if isinstance(pkl, dict) else load_pkl(pkl)
        if device is None:
            device = torch.cuda.current_device()
        self.device = torch.device('cuda', device if isinstance(device, int) else torch.device(device).index or 0)
        f32 = np.float32
        self.v_template = np.ascontiguousarray(dd['v_template'], dtype=f32)
        self.size = [self.v_template.shape[0], 3]
        self.num_betas = dd['shapedirs'].shape[-1]
        self.shapedirs = np.ascontiguousarray(np.reshape(dd['shapedirs'], [-1, self.num_betas]).T, dtype=f32)
        self.J_regressor = np.ascontiguousarray(np.asarray(dd['J_regressor'].T.todense()), dtype=f32)
        self.posedirs = np.ascontiguousarray(
            np.reshape(dd['posedirs'], [-1, dd['posedirs'].shape[-1]]).T, dtype=f32)
        self.parents = np.ascontiguousarray(dd['kintree_table'][0].astype(np.int32))
        self.weights = np.ascontiguousarray(dd['weights'], dtype=f32)
        jr = np.asarray(dd['cocoplus_regressor'].T.todense(), dtype=f32)
        if joint_type == 'lsp':
            jr = jr[:, :14]
        self.joint_regressor = np.ascontiguousarray(jr)
        if self.v_template.shape != (NV, 3) or self.num_betas != NB or self.posedirs.shape != (207, NV3):
            raise ValueError('not an SMPL-shaped model (6890 verts, 10 betas, 207 pose blend shapes)')
        self.NK = self.joint_regressor.shape[1]
        self._h = C.c_void_p()
        L = _lib.lib()

        def hp(a):
            return a.ctypes.data_as(C.c_void_p)

        check(L.smplb_model_create(C.byref(self._h), self.device.index, hp(self.v_template),
                                   hp(self.shapedirs), hp(self.posedirs), hp(self.J_regressor),
                                   hp(self.weights), hp(self.joint_regressor), self.NK,
                                   hp(self.parents)))
        self._ws = None
        self._last = None          # (workspace data_ptr, B, flags) of the last forward: what a backward may rely on
        self.n_forward_calls = 0   # smplb_forward launches made through this handle (tests count them)

    def close(self):
        if getattr(self, '_h', None) is not None and self._h.value:
            _lib.lib().smplb_model_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- workspace (caller-owned in the C ABI; cached here) ------------------------------------
    def workspace(self, B, flags):
        need = _lib.lib().smplb_workspace_bytes(self._h, B, flags)
        if self._ws is None or self._ws.numel() < need:
            self._ws = torch.empty(max(need, 256), dtype=torch.uint8, device=self.device)
        return self._ws, need

    def forward(self, beta, theta, cam_pos=None, cam_rot=None, focal=None, want_verts=True,
                want_kp=True, want_Rs=True, want_A=False, want_uv=True, flags=0, cam_cfg=None,
                out=None, ws=None, check_inputs=False):
        """One smplb_forward call.  Returns a dict of CUDA tensors.  check_inputs: also run smplb_forward_status
        (synchronises; raises SmplbError(E_NOT_FINITE) for a non-finite input or |beta| > 65504)."""
        dev = self.device
        beta = _dev_f32(beta, dev)
        theta = _dev_f32(theta, dev)
        B = beta.shape[0]
        if beta.shape != (B, NB) or theta.shape != (B, NJ * 3):
            raise ValueError('beta must be [B,10] and theta [B,72]; got %s %s' %
                             (tuple(beta.shape), tuple(theta.shape)))
        out = out if out is not None else {}

        def buf(name, shape, want):
            if not want:
                return None
            t = out.get(name)
            if t is None or tuple(t.shape) != tuple(shape):
                t = torch.empty(shape, dtype=torch.float32, device=dev)
                out[name] = t
            return t

        skin = not (flags & _lib.F_NO_SKIN)
        if skin and want_kp and not want_verts and (flags & _lib.F_SKIN_SIMT):
            flags |= _lib.F_SCRATCH_VERTS   # tensor-core path: keypoints come from the compact rows, no verts needed
        verts = buf('verts', (B, NV, 3), want_verts and skin)
        joints = buf('joints', (B, NJ, 3), True)
        kp = buf('kp', (B, self.NK, 3), want_kp and skin)
        Rs = buf('Rs', (B, NJ, 3, 3), want_Rs)
        A = buf('A', (B, NJ, 3, 4), want_A)
        cfg = dict(DEFAULT_CAM)
        cfg.update(cam_cfg or {})
        has_cam = cam_pos is not None
        if has_cam:
            cam_pos = _dev_f32(cam_pos, dev, (B, 3))
            cam_rot = _dev_f32(cam_rot, dev, (B, 3))
            focal = None if focal is None else _dev_f32(focal, dev, (B,))
        uv = buf('uv', (B, NJ, 2), has_cam and want_uv)
        px = buf('px', (B, NJ, 2), has_cam)
        io = ForwardIO(B, _ptr(beta), _ptr(theta), _ptr(verts), _ptr(joints), _ptr(kp), _ptr(Rs),
                       _ptr(A), _ptr(cam_pos) if has_cam else None,
                       _ptr(cam_rot) if has_cam else None, _ptr(focal) if has_cam else None,
                       _ptr(uv), _ptr(px), CamCfg(**cfg))
        if ws is None:
            ws, need = self.workspace(B, flags)
        with torch.cuda.device(dev):
            check(_lib.lib().smplb_forward(self._h, C.byref(io), _ptr(ws), ws.numel(), flags,
                                           _stream(dev)))
            self.n_forward_calls += 1
            self._last = (ws.data_ptr(), B, flags)
            if check_inputs:
                check(_lib.lib().smplb_forward_status(_ptr(ws), _stream(dev)))
        out['_keep'] = (beta, theta, cam_pos, cam_rot, focal)
        return out

    def new_workspace(self, B, flags):
        """A workspace of its own for one forward (+ backward) call: several forward calls can then stay alive until
        their backward runs (the reference's training step calls SMPL twice before differentiating,
        pose_model_3d.py:177-192)."""
        need = _lib.lib().smplb_workspace_bytes(self._h, B, flags)
        return torch.empty(max(need, 256), dtype=torch.uint8, device=self.device)

    def backward(self, beta, theta, g_verts=None, g_joints=None, g_kp=None, g_px=None,
                 cam_pos=None, cam_rot=None, focal=None, flags=0, cam_cfg=None, ws=None):
        """smplb_backward.  Vertex gradients (g_verts / g_kp) need the workspace of the matching
        forward(flags | F_KEEP_VPOSED): pass it as `ws`, or leave ws=None to use the handle's cached workspace, which
        is only accepted when the LAST forward through this handle filled it with the same batch and flags (a stale
        or foreign workspace would silently give wrong gradients)."""
        dev = self.device
        beta = _dev_f32(beta, dev)
        theta = _dev_f32(theta, dev)
        B = beta.shape[0]
        g_beta = torch.empty((B, NB), dtype=torch.float32, device=dev)
        g_theta = torch.empty((B, NJ * 3), dtype=torch.float32, device=dev)
        cfg = dict(DEFAULT_CAM)
        cfg.update(cam_cfg or {})

        def opt(x, shape):
            return None if x is None else _dev_f32(x, dev, shape)

        g_verts = opt(g_verts, (B, NV, 3))
        g_joints = opt(g_joints, (B, NJ, 3))
        g_kp = opt(g_kp, (B, self.NK, 3))
        g_px = opt(g_px, (B, NJ, 2))
        cam_pos = opt(cam_pos, (B, 3))
        cam_rot = opt(cam_rot, (B, 3))
        focal = opt(focal, (B,))
        io = BackwardIO(B, _ptr(beta), _ptr(theta), _ptr(g_verts), _ptr(g_joints), _ptr(g_kp),
                        _ptr(g_px), _ptr(cam_pos), _ptr(cam_rot), _ptr(focal), CamCfg(**cfg),
                        _ptr(g_beta), _ptr(g_theta))
        if g_verts is not None or g_kp is not None:   # vertex path: needs the forward's workspace
            flags |= _lib.F_KEEP_VPOSED
            if ws is None:
                ws = self._ws
                layout = _lib.F_KEEP_VPOSED | _lib.F_NO_SKIN | _lib.F_SCRATCH_VERTS | _lib.F_FUSED   # what shapes the workspace
                if ws is None or self._last is None or self._last[:2] != (ws.data_ptr(), B) or \
                        (self._last[2] & layout) != (flags & layout):
                    raise RuntimeError(
                        'Model.backward: the cached workspace does not hold the forward of this call (last forward: %r, '
                        'wanted B=%d flags=%d incl. F_KEEP_VPOSED); run forward(flags | F_KEEP_VPOSED) right before, or '
                        'give each call its own workspace (new_workspace / ws=)' % (self._last, B, flags))
            if ws.numel() < _lib.lib().smplb_workspace_bytes(self._h, B, flags):
                raise RuntimeError('Model.backward: workspace too small for B=%d flags=%d' % (B, flags))
            wsp, wsn = _ptr(ws), ws.numel()
        else:
            wsp, wsn = None, 0
        with torch.cuda.device(dev):
            check(_lib.lib().smplb_backward(self._h, C.byref(io), wsp, wsn, flags, _stream(dev)))
        return g_beta, g_theta


# -- stand-alone operators -------------------------------------------------------------------------

def _cuda_dev(x):
    if torch.is_tensor(x) and x.is_cuda:
        return x.device
    return torch.device('cuda', torch.cuda.current_device())


def rodrigues(rvecs, variant):
    """variant 0: batch_rodrigues (batch_lbs.py:38-57); 1: rodrigues_batch (project.py:51-81)."""
    dev = _cuda_dev(rvecs)
    r = _dev_f32(rvecs, dev)
    N = r.shape[0]
    out = torch.empty((N, 3, 3), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        check(_lib.lib().smplb_rodrigues(N, _ptr(r), _ptr(out), variant, _stream(dev)))
    return out


def global_rigid(Rs, Js, parents, rotate_base=False):
    """batch_global_rigid_transformation (batch_lbs.py:84-144)."""
    dev = _cuda_dev(Rs)
    Rs = _dev_f32(Rs, dev)
    N = Rs.shape[0]
    Rs = Rs.reshape(N, NJ, 3, 3)
    Js = _dev_f32(Js, dev, (N, NJ, 3))
    par = np.ascontiguousarray(np.asarray(parents).astype(np.int32))
    newJ = torch.empty((N, NJ, 3), dtype=torch.float32, device=dev)
    A = torch.empty((N, NJ, 4, 4), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        check(_lib.lib().smplb_global_rigid(N, _ptr(Rs), _ptr(Js), par.ctypes.data_as(C.c_void_p),
                                            int(bool(rotate_base)), _ptr(newJ), _ptr(A), _stream(dev)))
    return newJ, A


def project(points, cams_pos, cams_angles, focal_lens, P=1, cam_tiled=True):
    """project (project.py:7-48)."""
    dev = _cuda_dev(points)
    pts = _dev_f32(points, dev)
    N = pts.shape[0]
    cp = _dev_f32(cams_pos, dev)
    ca = _dev_f32(cams_angles, dev)
    fl = _dev_f32(focal_lens, dev)
    nc = N if cam_tiled else N // max(P, 1)
    if pts.shape != (N, 3) or cp.shape != (nc, 3) or ca.shape != (nc, 3) or fl.shape != (nc,):
        raise ValueError('project: inconsistent shapes')
    uv = torch.empty((N, 2), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        check(_lib.lib().smplb_project(N, P, _ptr(pts), _ptr(cp), _ptr(ca), _ptr(fl),
                                       int(bool(cam_tiled)), _ptr(uv), _stream(dev)))
    return uv


def project_backward(points, cams_pos, cams_angles, focal_lens, g_uv, P=1, cam_tiled=True, want_focal=False):
    """Adjoint of project(): (g_pts, g_cam_pos, g_cam_angles) and, with want_focal, g_focal_lens as a 4th value."""
    dev = _cuda_dev(points)
    pts = _dev_f32(points, dev)
    N = pts.shape[0]
    cp = _dev_f32(cams_pos, dev)
    ca = _dev_f32(cams_angles, dev)
    fl = _dev_f32(focal_lens, dev)
    g = _dev_f32(g_uv, dev, (N, 2))
    g_pts = torch.empty((N, 3), dtype=torch.float32, device=dev)
    g_cp = torch.empty_like(cp)
    g_ca = torch.empty_like(ca)
    g_fl = torch.empty_like(fl) if want_focal else None
    with torch.cuda.device(dev):
        check(_lib.lib().smplb_project_backward(N, P, _ptr(pts), _ptr(cp), _ptr(ca), _ptr(fl),
                                                int(bool(cam_tiled)), _ptr(g), _ptr(g_pts),
                                                _ptr(g_cp), _ptr(g_ca), _ptr(g_fl), _stream(dev)))
    return (g_pts, g_cp, g_ca, g_fl) if want_focal else (g_pts, g_cp, g_ca)


def orth_proj(X, camera):
    """batch_orth_proj_idrot (projection.py:10-26)."""
    dev = _cuda_dev(X)
    X = _dev_f32(X, dev)
    N, P = X.shape[0], X.shape[1]
    cam = _dev_f32(camera, dev).reshape(N, 3)
    out = torch.empty((N, P, 2), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        check(_lib.lib().smplb_orth_proj(N, P, _ptr(X), _ptr(cam), _ptr(out), _stream(dev)))
    return out


def reproj_loss(px_pred, px_gt, ss=0.024, scale=None, cap=1.0 / 24):
    """pose_model_3d.py:233-237 -> (uncapped, capped) as a 2-element CUDA tensor."""
    dev = _cuda_dev(px_pred)
    a = _dev_f32(px_pred, dev)
    g = _dev_f32(px_gt, dev, a.shape)
    B = a.shape[0]
    cfg = LossCfg(CamCfg(**DEFAULT_CAM), ss if scale is None else scale, cap if cap else 0.0, 2.0, 1.0)
    out = torch.empty(2, dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        check(_lib.lib().smplb_reproj_loss(B, _ptr(a), _ptr(g), C.byref(cfg), _ptr(out), _stream(dev)))
    return out


def mse_grad(pred, target, weight=1.0, n_total=None, want_grad=True, loss_out=None, accumulate=False, g_out=None,
             scratch=None):
    """tf.losses.mean_squared_error with a scalar weight (pose_model_3d.py:183-192): returns (loss[1] CUDA tensor,
    gradient w.r.t. pred or None).  accumulate adds into loss_out."""
    dev = _cuda_dev(pred)
    a = _dev_f32(pred, dev)
    b = _dev_f32(target, dev, a.shape)
    n = a.numel()
    if loss_out is None:
        loss_out = torch.zeros(1, dtype=torch.float32, device=dev)
    if want_grad and g_out is None:
        g_out = torch.empty_like(a)
    if scratch is None:
        scratch = torch.empty(int(_lib.lib().smplb_mse_scratch_bytes()), dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        check(_lib.lib().smplb_mse_grad(n, _ptr(a), _ptr(b), float(weight), int(n_total if n_total is not None else n),
                                        _ptr(g_out) if want_grad else None, _ptr(loss_out), int(bool(accumulate)),
                                        _ptr(scratch), _stream(dev)))
    return loss_out, (g_out if want_grad else None)


def adam_step(x, g, m, v, t, lr=2e-4, beta1=0.9, beta2=0.999, eps=1e-8):
    """In-place TF1 Adam (pose_model_3d.py:393)."""
    dev = x.device
    assert x.is_cuda and x.is_contiguous() and x.dtype == torch.float32
    with torch.cuda.device(dev):
        check(_lib.lib().smplb_adam_step(x.numel(), _ptr(x), _ptr(g), _ptr(m), _ptr(v), int(t),
                                         lr, beta1, beta2, eps, _stream(dev)))
    return x


def adam_tick(step_dev, lr_t_dev, lr=2e-4, beta1=0.9, beta2=0.999):
    """t = ++step_dev[0] (int32 device counter), lr_t_dev[0] = lr sqrt(1-b2^t)/(1-b1^t): once per optimiser step;
    graph-capturable (the counter advances on every replay)."""
    dev = step_dev.device
    assert step_dev.is_cuda and step_dev.dtype == torch.int32 and lr_t_dev.dtype == torch.float32
    with torch.cuda.device(dev):
        check(_lib.lib().smplb_adam_tick(_ptr(step_dev), lr, beta1, beta2, _ptr(lr_t_dev), _stream(dev)))
    return lr_t_dev


def adam_step_dev(x, g, m, v, lr_t_dev, beta1=0.9, beta2=0.999, eps=1e-8):
    """In-place TF1 Adam (pose_model_3d.py:393) with lr_t read from device memory (see adam_tick)."""
    dev = x.device
    assert x.is_cuda and x.is_contiguous() and x.dtype == torch.float32
    with torch.cuda.device(dev):
        check(_lib.lib().smplb_adam_step_dev(x.numel(), _ptr(x), _ptr(g), _ptr(m), _ptr(v), _ptr(lr_t_dev),
                                             beta1, beta2, eps, _stream(dev)))
    return x


# -- pose pre-processing and the camera losses ------------------------------------

def add_axis_angle_rotations(rv1, rv2):
    """deps/pose_3d/utils.py:170-199: [N,3], [N,3] -> [N,3]."""
    dev = _cuda_dev(rv1)
    rv1, rv2 = _dev_f32(rv1, dev), _dev_f32(rv2, dev)
    out = torch.empty_like(rv1)
    with torch.cuda.device(dev):
        check(_lib.lib().smplb_add_axis_angle(rv1.shape[0], _ptr(rv1), _ptr(rv2), _ptr(out), _stream(dev)))
    return out


def rotate_global_pose(thetas, zrot):
    """deps/pose_3d/utils.py:152-167: thetas [B,72], zrot [B] -> [B,72]."""
    dev = _cuda_dev(thetas)
    thetas, zrot = _dev_f32(thetas, dev), _dev_f32(zrot, dev)
    out = torch.empty_like(thetas)
    with torch.cuda.device(dev):
        check(_lib.lib().smplb_rotate_global_pose(thetas.shape[0], _ptr(thetas), _ptr(zrot), _ptr(out), _stream(dev)))
    return out


def cam_losses(gt_joints3d, gt_joints2d, cam_pos, cam_rot, cam_f, f_raw=None, angle_scale=10.0, cam_cfg=None,
               want_grad=True):
    """pose_model_3d.py:256-292.  Returns (loss[4] = angle, neg, rot, sum; grads dict or None)."""
    dev = _cuda_dev(gt_joints3d)
    B = gt_joints3d.shape[0]
    j3, j2 = _dev_f32(gt_joints3d, dev, (B, NJ, 3)), _dev_f32(gt_joints2d, dev, (B, NJ, 2))
    cp, cr, cf = _dev_f32(cam_pos, dev, (B, 3)), _dev_f32(cam_rot, dev, (B, 3)), _dev_f32(cam_f, dev, (B,))
    fr = None if f_raw is None else _dev_f32(f_raw, dev, (B,))
    cfg = dict(DEFAULT_CAM)
    cfg.update(cam_cfg or {})
    loss = torch.empty(4, dtype=torch.float32, device=dev)
    g = None
    if want_grad:
        g = {'cam_pos': torch.empty_like(cp), 'cam_rot': torch.empty_like(cr), 'cam_f': torch.empty_like(cf),
             'f_raw': None if fr is None else torch.empty_like(fr)}
    scratch = torch.empty(max(int(_lib.lib().smplb_cam_losses_scratch_bytes(B)), 16), dtype=torch.uint8, device=dev)
    cc = CamCfg(**cfg)
    with torch.cuda.device(dev):
        check(_lib.lib().smplb_cam_losses(B, _ptr(j3), _ptr(j2), _ptr(cp), _ptr(cr), _ptr(cf), _ptr(fr), C.byref(cc),
                                          float(angle_scale), _ptr(loss), _ptr(g['cam_pos']) if g else None,
                                          _ptr(g['cam_rot']) if g else None, _ptr(g['cam_f']) if g else None,
                                          _ptr(g['f_raw']) if g and fr is not None else None, _ptr(scratch),
                                          _stream(dev)))
    return loss, g
