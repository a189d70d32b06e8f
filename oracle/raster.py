"""TEST INFRASTRUCTURE ONLY (oracle): ctypes access to the rasteriser oracles.

  ref  = the reference's own C++ (rasterize_triangles_impl.cc compiled from /root/reference into
         oracle/_ref/librasterize_ref.so by oracle/Makefile) -- forward only
  port = oracle/raster_port.c, our plain-C restatement (forward + the gradient op)
Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may import this module."""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
_port = _ref = None


def _fp(a):
    return a.ctypes.data_as(C.c_void_p)


def port():
    global _port
    if _port is None:
        path = os.path.join(HERE, 'libraster_port.so')
        if not os.path.exists(path):
            raise RuntimeError('oracle/libraster_port.so missing: run `make -C oracle port` (or __graft_entry__.build())')
        _port = C.CDLL(path)
    return _port


def ref_available():
    return os.path.exists(os.path.join(HERE, '_ref', 'librasterize_ref.so'))


def ref():
    global _ref
    if _ref is None:
        _ref = C.CDLL(os.path.join(HERE, '_ref', 'librasterize_ref.so'))
    return _ref


def _forward(fn, ndc, triangles, W, H):
    ndc = np.ascontiguousarray(ndc, dtype=np.float32)
    tri = np.ascontiguousarray(triangles, dtype=np.int32)
    ids = np.empty((H, W), np.int32)
    bary = np.empty((H, W, 3), np.float32)
    z = np.empty((H, W), np.float32)
    fn(_fp(ndc), _fp(tri), C.c_int32(tri.shape[0]), C.c_int32(W), C.c_int32(H), _fp(ids), _fp(bary), _fp(z))
    return bary, ids, z


def forward_port(ndc, triangles, W, H):
    """One image: ndc [V,3] NDC vertices, triangles [T,3] -> (bary [H,W,3], ids [H,W], z [H,W])."""
    return _forward(port().raster_port_forward, ndc, triangles, W, H)


def forward_ref(ndc, triangles, W, H):
    return _forward(ref().raster_ref_forward, ndc, triangles, W, H)


def backward_port(ndc, triangles, bary, ids, g_bary, W, H):
    ndc = np.ascontiguousarray(ndc, dtype=np.float32)
    tri = np.ascontiguousarray(triangles, dtype=np.int32)
    bary = np.ascontiguousarray(bary, dtype=np.float32)
    ids = np.ascontiguousarray(ids, dtype=np.int32)
    g_bary = np.ascontiguousarray(g_bary, dtype=np.float32)
    g = np.empty_like(ndc)
    port().raster_port_backward(_fp(ndc), C.c_int32(ndc.shape[0]), _fp(tri), _fp(bary), _fp(ids), _fp(g_bary),
                                C.c_int32(W), C.c_int32(H), _fp(g))
    return g
