"""Rasteriser bindings: smplb_mesh handle + batched rasterise / gradient (include/smplb200.h).
torch tensors are carriers only; every pixel is computed by the kernels in csrc/raster.cu."""
import ctypes as C

import numpy as np
import torch

from . import _lib
from ._lib import check


def _ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


class Mesh:
    """Triangle list [T,3] (int32, clockwise = front facing) of a mesh with V vertices, resident on `device`."""

    def __init__(self, triangles, vertex_count, device=None):
        self._h = None
        if device is None:
            device = torch.cuda.current_device()
        self.device = torch.device('cuda', device) if isinstance(device, int) else torch.device(device)
        tri = np.ascontiguousarray(np.asarray(triangles.cpu() if torch.is_tensor(triangles) else triangles),
                                   dtype=np.int32).reshape(-1, 3)
        self.T, self.V = int(tri.shape[0]), int(vertex_count)
        h = C.c_void_p()
        check(_lib.lib().smplb_mesh_create(C.byref(h), self.device.index or 0, tri.ctypes.data_as(C.c_void_p),
                                           self.T, self.V))
        self._h = h
        self.triangles = torch.from_numpy(tri.astype(np.int64)).to(self.device)
        self._ws = None

    def close(self):
        if self._h is not None:
            _lib.lib().smplb_mesh_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _workspace(self, B, W, H):
        need = _lib.lib().smplb_rasterize_workspace_bytes(self._h, B, W, H)
        if self._ws is None or self._ws.numel() < need:
            self._ws = torch.empty(max(need, 256), dtype=torch.uint8, device=self.device)
        return self._ws

    def rasterize(self, ndc, W, H):
        """ndc [B,V,3] -> (bary [B,H,W,3] f32, ids [B,H,W] i32, z [B,H,W] f32)."""
        ndc = ndc.to(self.device, torch.float32).contiguous()
        if ndc.dim() != 3 or ndc.shape[1] != self.V or ndc.shape[2] != 3:
            raise ValueError('ndc must be [B,%d,3]; got %s' % (self.V, tuple(ndc.shape)))
        B = ndc.shape[0]
        bary = torch.empty((B, H, W, 3), dtype=torch.float32, device=self.device)
        ids = torch.empty((B, H, W), dtype=torch.int32, device=self.device)
        z = torch.empty((B, H, W), dtype=torch.float32, device=self.device)
        ws = self._workspace(B, W, H)
        with torch.cuda.device(self.device):
            st = C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
            check(_lib.lib().smplb_rasterize(self._h, B, _ptr(ndc), W, H, _ptr(bary), _ptr(ids), _ptr(z), _ptr(ws),
                                             ws.numel(), st))
        return bary, ids, z

    def normals(self, verts, out=None):
        """normals_from_mesh (deps/pose_3d/utils.py:92-127): verts [B,V,3] -> [B,V,3]."""
        verts = verts.to(self.device, torch.float32).contiguous()
        if verts.dim() != 3 or verts.shape[1] != self.V or verts.shape[2] != 3:
            raise ValueError('verts must be [B,%d,3]; got %s' % (self.V, tuple(verts.shape)))
        if out is None:
            out = torch.empty_like(verts)
        with torch.cuda.device(self.device):
            st = C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
            check(_lib.lib().smplb_mesh_normals(self._h, verts.shape[0], _ptr(verts), _ptr(out), st))
        return out

    def render_diffuse(self, verts, normals, clip_transforms, light_positions, light_intensities, W, H,
                       diffuse=None, ambient=None):
        """smplb_render_diffuse: one fused pass verts/normals -> shaded RGBA [B,H,W,4] (no specular term)."""
        dev = self.device

        def f32(x, shape):
            x = x.to(dev, torch.float32).contiguous()
            if tuple(x.shape) != tuple(shape):
                raise ValueError('expected shape %s, got %s' % (tuple(shape), tuple(x.shape)))
            return x

        B = verts.shape[0]
        verts, normals = f32(verts, (B, self.V, 3)), f32(normals, (B, self.V, 3))
        diffuse = None if diffuse is None else f32(diffuse, (B, self.V, 3))
        clip_transforms = f32(clip_transforms, (B, 4, 4))
        L = light_positions.shape[1]
        lp, li = f32(light_positions, (B, L, 3)), f32(light_intensities, (B, L, 3))
        ambient = None if ambient is None else f32(ambient, (B, 3))
        rgba = torch.empty((B, H, W, 4), dtype=torch.float32, device=dev)
        need = _lib.lib().smplb_render_workspace_bytes(self._h, B, W, H)
        if self._ws is None or self._ws.numel() < need:
            self._ws = torch.empty(max(need, 256), dtype=torch.uint8, device=dev)
        with torch.cuda.device(dev):
            st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
            check(_lib.lib().smplb_render_diffuse(self._h, B, _ptr(verts), _ptr(normals), _ptr(diffuse),
                                                  _ptr(clip_transforms), _ptr(lp), _ptr(li), L, _ptr(ambient), W, H,
                                                  _ptr(rgba), _ptr(self._ws), self._ws.numel(), st))
        return rgba

    def rasterize_backward(self, ndc, bary, ids, g_bary, W, H):
        ndc = ndc.to(self.device, torch.float32).contiguous()
        B = ndc.shape[0]
        g = torch.empty_like(ndc)
        ws = self._workspace(B, W, H)
        with torch.cuda.device(self.device):
            st = C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
            check(_lib.lib().smplb_rasterize_backward(self._h, B, _ptr(ndc), _ptr(bary.contiguous()),
                                                      _ptr(ids.contiguous()), _ptr(g_bary.contiguous().float()), W, H,
                                                      _ptr(g), _ptr(ws), ws.numel(), st))
        return g


class _Rasterize(torch.autograd.Function):
    """RasterizeTriangles with the registered gradient (rasterize_triangles.py:181-188): gradients flow through the
    barycentrics only; ids and z are not differentiable."""

    @staticmethod
    def forward(ctx, ndc, mesh, W, H):
        bary, ids, z = mesh.rasterize(ndc, W, H)
        ctx.mesh, ctx.W, ctx.H = mesh, W, H
        ctx.save_for_backward(ndc, bary, ids)
        ctx.mark_non_differentiable(ids, z)
        return bary, ids, z

    @staticmethod
    def backward(ctx, g_bary, g_ids, g_z):
        ndc, bary, ids = ctx.saved_tensors
        return ctx.mesh.rasterize_backward(ndc, bary, ids, g_bary, ctx.W, ctx.H), None, None, None


def rasterize_ndc(ndc, mesh, W, H):
    return _Rasterize.apply(ndc, mesh, W, H)


def _stream(dev):
    return C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)


class _ClipTransform(torch.autograd.Function):
    """rasterize_triangles.py:73-86 (csrc/interp.cu): verts [B,V,3], clip matrices [B,4,4] -> ndc [B,V,3], clip_w [B,V]."""

    @staticmethod
    def forward(ctx, verts, clip_tf):
        dev = verts.device
        verts, clip_tf = verts.to(torch.float32).contiguous(), clip_tf.to(dev, torch.float32).contiguous()
        B, V = verts.shape[0], verts.shape[1]
        ndc = torch.empty((B, V, 3), dtype=torch.float32, device=dev)
        cw = torch.empty((B, V), dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            check(_lib.lib().smplb_clip_transform(B, V, _ptr(verts), _ptr(clip_tf), _ptr(ndc), _ptr(cw), _stream(dev)))
        ctx.save_for_backward(verts, clip_tf)
        return ndc, cw

    @staticmethod
    def backward(ctx, g_ndc, g_cw):
        verts, clip_tf = ctx.saved_tensors
        dev = verts.device
        B, V = verts.shape[0], verts.shape[1]
        g_ndc = torch.zeros_like(verts) if g_ndc is None else g_ndc.to(torch.float32).contiguous()
        g_cw = None if g_cw is None else g_cw.to(torch.float32).contiguous()
        g = torch.empty_like(verts)
        with torch.cuda.device(dev):
            check(_lib.lib().smplb_clip_transform_backward(B, V, _ptr(verts), _ptr(clip_tf), _ptr(g_ndc), _ptr(g_cw), _ptr(g),
                                                           _stream(dev)))
        return g, None


class _Interpolate(torch.autograd.Function):
    """rasterize_triangles.py:119-164 (csrc/interp.cu): attributes [B,V,C], clip_w [B,V], bary [B,H,W,3], ids [B,H,W],
    background [C] -> images [B,H,W,C]; differentiable wrt attributes, clip_w and bary."""

    @staticmethod
    def forward(ctx, attr, clip_w, bary, ids, bg, mesh):
        dev = attr.device
        f = lambda x: x.to(dev, torch.float32).contiguous()
        attr, clip_w, bary, bg = f(attr), f(clip_w), f(bary), f(bg)
        ids = ids.to(dev, torch.int32).contiguous()
        B, H, W = ids.shape
        Cn = attr.shape[2]
        out = torch.empty((B, H, W, Cn), dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            check(_lib.lib().smplb_interpolate(mesh._h, B, Cn, _ptr(attr), _ptr(clip_w), _ptr(bary), _ptr(ids), _ptr(bg), W, H,
                                               _ptr(out), _stream(dev)))
        ctx.mesh = mesh
        ctx.save_for_backward(attr, clip_w, bary, ids, bg)
        return out

    @staticmethod
    def backward(ctx, g_out):
        attr, clip_w, bary, ids, bg = ctx.saved_tensors
        dev = attr.device
        B, H, W = ids.shape
        Cn = attr.shape[2]
        g_out = g_out.to(torch.float32).contiguous()
        g_attr, g_cw, g_bary = torch.empty_like(attr), torch.empty_like(clip_w), torch.empty_like(bary)
        with torch.cuda.device(dev):
            check(_lib.lib().smplb_interpolate_backward(ctx.mesh._h, B, Cn, _ptr(attr), _ptr(clip_w), _ptr(bary), _ptr(ids),
                                                        _ptr(bg), W, H, _ptr(g_out), _ptr(g_attr), _ptr(g_cw), _ptr(g_bary),
                                                        _stream(dev)))
        return g_attr, g_cw, g_bary, None, None, None


def clip_transform(verts, clip_tf):
    return _ClipTransform.apply(verts, clip_tf)


def interpolate(attr, clip_w, bary, ids, background, mesh):
    return _Interpolate.apply(attr, clip_w, bary, ids, background, mesh)
