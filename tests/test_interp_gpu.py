"""Clip transform and attribute interpolation kernels with their adjoints (csrc/interp.cu) against the float64 torch
restatement of deps/tf_mesh_renderer/rasterize_triangles.py:73-164 with autograd (oracle/ref_torch.py).
Tolerances: values 1e-5 max-norm relative, gradients 1e-4."""
import numpy as np
import pytest
import torch

from conftest import maxrel

pytestmark = pytest.mark.gpu


def test_clip_transform_forward_and_adjoint():
    from oracle import ref_torch as rt
    from smplb200.raster import clip_transform
    g = torch.Generator().manual_seed(2)
    B, V = 3, 257
    v64 = (torch.randn(B, V, 3, generator=g, dtype=torch.float64)).requires_grad_(True)
    M64 = torch.randn(B, 4, 4, generator=g, dtype=torch.float64)
    M64[:, 3, :3] *= 0.1       # keep the clip w away from zero: the divide amplifies fp32 rounding there
    M64[:, 3, 3] = 3.0
    w1 = torch.randn(B, V, 3, generator=g, dtype=torch.float64)
    w2 = torch.randn(B, V, generator=g, dtype=torch.float64)
    ndc64, cw64 = rt.clip_transform(v64, M64)
    ((ndc64 * w1).sum() + (cw64 * w2).sum()).backward()
    v = v64.detach().float().cuda().requires_grad_(True)
    ndc, cw = clip_transform(v, M64.float().cuda())
    assert maxrel(ndc.detach().cpu().double().numpy(), ndc64.detach().numpy()) <= 1e-5
    assert maxrel(cw.detach().cpu().double().numpy(), cw64.detach().numpy()) <= 1e-5
    ((ndc * w1.float().cuda()).sum() + (cw * w2.float().cuda()).sum()).backward()
    assert maxrel(v.grad.cpu().double().numpy(), v64.grad.numpy()) <= 1e-4


def test_interpolate_forward_and_adjoint():
    from oracle import ref_torch as rt
    from smplb200 import synthetic as syn
    from smplb200.raster import Mesh, interpolate
    vv, tri = syn.uv_sphere_mesh(12, 10)
    V, T = vv.shape[0], tri.shape[0]
    g = torch.Generator().manual_seed(4)
    B, H, W, C = 2, 9, 11, 7
    attr64 = torch.randn(B, V, C, generator=g, dtype=torch.float64).requires_grad_(True)
    cw64 = (torch.rand(B, V, generator=g, dtype=torch.float64) * 2.0 + 0.5).requires_grad_(True)
    bary = torch.rand(B, H, W, 3, generator=g, dtype=torch.float64)
    bary = bary / bary.sum(3, keepdim=True)
    bgmask = torch.rand(B, H, W, generator=g) < 0.3            # background pixels: zero barycentrics, id 0
    bary[bgmask] = 0.0
    bary64 = bary.clone().requires_grad_(True)
    ids = torch.randint(0, T, (B, H, W), generator=g)
    ids[bgmask] = 0
    bg = torch.linspace(-1.0, 0.5, C, dtype=torch.float64)
    wgt = torch.randn(B, H, W, C, generator=g, dtype=torch.float64)
    want = rt.interpolate(attr64, cw64, torch.from_numpy(tri.astype(np.int64)), bary64, ids, bg)
    (want * wgt).sum().backward()

    mesh = Mesh(tri, V, 0)
    c = lambda x: x.detach().float().cuda().requires_grad_(True)
    attr, cw, b32 = c(attr64), c(cw64), c(bary64)
    got = interpolate(attr, cw, b32, ids.int().cuda(), bg.float().cuda(), mesh)
    assert maxrel(got.detach().cpu().double().numpy(), want.detach().numpy()) <= 1e-5
    (got * wgt.float().cuda()).sum().backward()
    assert maxrel(attr.grad.cpu().double().numpy(), attr64.grad.numpy()) <= 1e-4
    assert maxrel(cw.grad.cpu().double().numpy(), cw64.grad.numpy()) <= 1e-4
    # background pixels carry 1 / 1e-6 in d/d bary (as the reference graph does): compare covered and background apart
    gb, gb64 = b32.grad.cpu().double(), bary64.grad
    assert maxrel(gb[~bgmask].numpy(), gb64[~bgmask].numpy()) <= 1e-4
    assert maxrel(gb[bgmask].numpy(), gb64[bgmask].numpy()) <= 1e-4


def test_new_render_entry_points_reject_bad_arguments():
    """Error behaviour of the round-2 renderer entry points through the raw C ABI: a status code and a message, never a
    launch on bad arguments."""
    import ctypes as C
    from smplb200 import _lib, synthetic as syn
    from smplb200.raster import Mesh
    L = _lib.lib()
    vv, tri = syn.uv_sphere_mesh(8, 6)
    mesh = Mesh(tri, vv.shape[0], 0)
    buf = torch.zeros(4096, device='cuda')
    p = C.c_void_p(buf.data_ptr())
    assert L.smplb_clip_transform(-1, 4, p, p, p, p, None) == _lib.E_BAD_SHAPE
    assert L.smplb_clip_transform(1, 4, None, p, p, p, None) == _lib.E_INVALID_ARG
    assert L.smplb_clip_transform(0, 4, None, None, None, None, None) == 0          # empty batch: nothing to do
    assert L.smplb_clip_transform_backward(1, 4, p, p, None, None, p, None) == _lib.E_INVALID_ARG
    assert L.smplb_interpolate(None, 1, 3, p, p, p, p, p, 4, 4, p, None) == _lib.E_INVALID_ARG
    assert L.smplb_interpolate(mesh._h, 1, 0, p, p, p, p, p, 4, 4, p, None) == _lib.E_BAD_SHAPE
    assert L.smplb_interpolate(mesh._h, 1, 3, p, p, None, p, p, 4, 4, p, None) == _lib.E_INVALID_ARG
    assert L.smplb_interpolate_backward(mesh._h, 1, 3, p, p, p, p, p, 4, 4, p, p, None, p, None) == _lib.E_INVALID_ARG
    io = _lib.PhongIO(1, 4, 4, 9, 9, 0, 3, 6, -1, -1, 0, p, None, None, p, p, None, None)          # 9 lights > 8
    assert L.smplb_phong_shade(C.byref(io), p, p, 1 << 14, None) == _lib.E_BAD_SHAPE
    io = _lib.PhongIO(1, 4, 4, 9, 1, 0, 3, 8, -1, -1, 0, p, None, None, p, p, None, None)          # diffuse offset outside C
    assert L.smplb_phong_shade(C.byref(io), p, p, 1 << 14, None) == _lib.E_BAD_SHAPE
    io = _lib.PhongIO(1, 4, 4, 12, 1, 0, 3, 6, 9, -1, 0, p, None, None, p, p, None, None)          # specular without camera
    assert L.smplb_phong_shade(C.byref(io), p, p, 1 << 14, None) == _lib.E_INVALID_ARG
    io = _lib.PhongIO(1, 4, 4, 9, 1, 0, 3, 6, -1, -1, 0, p, None, None, p, p, None, None)
    assert L.smplb_phong_shade(C.byref(io), p, p, 8, None) == _lib.E_INVALID_ARG                   # scratch too small
    assert b'scratch' in L.smplb_last_error()
    need = L.smplb_phong_scratch_bytes(1, 4, 4, 1)
    assert L.smplb_phong_shade(C.byref(io), p, p, need, None) == 0
    torch.cuda.synchronize()
