"""Phong shading kernels (csrc/shade.cu): smplb_phong_shade / smplb_phong_shade_backward behind a torch.autograd.Function.
torch tensors are carriers only; the lit image and its gradient are computed by the kernels."""
import ctypes as C

import torch

from . import _lib
from ._lib import check


def _ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


class _Phong(torch.autograd.Function):
    """attr [B,H,W,C] (+ optional alphas [B,H,W]) -> RGBA [B,H,W,4], rows reversed (mesh_renderer.py:23-162).
    Differentiable wrt attr and alphas; lights, camera, ambient and per-image shininess are constants."""

    @staticmethod
    def forward(ctx, attr, alphas, offs, normalize, shininess, light_pos, light_int, camera, ambient):
        dev = attr.device
        B, H, W, Cn = attr.shape
        L = light_pos.shape[1]
        f = lambda x: None if x is None else x.to(dev, torch.float32).contiguous()
        attr, alphas, shininess, light_pos, light_int, camera, ambient = (
            f(attr), f(alphas), f(shininess), f(light_pos), f(light_int), f(camera), f(ambient))
        io = _lib.PhongIO(B, H, W, Cn, L, offs[0], offs[1], offs[2], offs[3], offs[4], 1 if normalize else 0,
                          _ptr(attr), _ptr(alphas), _ptr(shininess), _ptr(light_pos), _ptr(light_int), _ptr(camera),
                          _ptr(ambient))
        need = _lib.lib().smplb_phong_scratch_bytes(B, H, W, L)
        scratch = torch.empty(max(need, 256), dtype=torch.uint8, device=dev)
        rgba = torch.empty((B, H, W, 4), dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
            check(_lib.lib().smplb_phong_shade(C.byref(io), _ptr(rgba), _ptr(scratch), scratch.numel(), st))
        ctx.keep = (attr, alphas, shininess, light_pos, light_int, camera, ambient, scratch)   # the io holds raw pointers
        ctx.io = io
        return rgba

    @staticmethod
    def backward(ctx, g_rgba):
        attr, alphas = ctx.keep[0], ctx.keep[1]
        scratch = ctx.keep[-1]
        dev = attr.device
        g_rgba = g_rgba.to(dev, torch.float32).contiguous()
        g_attr = torch.empty_like(attr)
        g_alphas = torch.empty_like(alphas) if alphas is not None else None
        with torch.cuda.device(dev):
            st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
            check(_lib.lib().smplb_phong_shade_backward(C.byref(ctx.io), _ptr(g_rgba), _ptr(g_attr), _ptr(g_alphas),
                                                        _ptr(scratch), scratch.numel(), st))
        return g_attr, g_alphas, None, None, None, None, None, None, None


def phong_shade(attr, light_positions, light_intensities, off_normal=0, off_position=3, off_diffuse=6, off_specular=-1,
                off_shininess=-1, alphas=None, normalize_normals=False, shininess=None, camera_position=None,
                ambient_color=None):
    """Lit RGBA of an interleaved attribute image (see smplb_phong_io in include/smplb200.h)."""
    offs = (off_normal, off_position, off_diffuse, off_specular, off_shininess)
    return _Phong.apply(attr, alphas, offs, normalize_normals, shininess, light_positions, light_intensities,
                        camera_position, ambient_color)
