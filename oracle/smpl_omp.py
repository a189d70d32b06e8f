"""ctypes front of oracle/smpl_omp.c: the tight C + OpenMP CPU version of the hot path.
TEST / BENCH INFRASTRUCTURE: a CPU number timed beside the GPU path and checked against the float64 oracle; the product
never imports it."""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def lib():
    global _LIB
    if _LIB is None:
        so = os.path.join(_HERE, 'libsmpl_omp.so')
        if not os.path.exists(so):
            subprocess.check_call(['make', '-C', _HERE, 'libsmpl_omp.so'])
        L = ctypes.CDLL(so)
        fp, ip = ctypes.POINTER(ctypes.c_float), ctypes.POINTER(ctypes.c_int32)
        L.smpl_omp_create.restype = ctypes.c_void_p
        L.smpl_omp_create.argtypes = [fp, fp, fp, fp, fp, fp, ctypes.c_int, ip]
        L.smpl_omp_free.argtypes = [ctypes.c_void_p]
        L.smpl_omp_forward.restype = ctypes.c_int
        L.smpl_omp_forward.argtypes = [ctypes.c_void_p, ctypes.c_int, fp, fp, fp, fp, fp, ctypes.c_float, ctypes.c_float,
                                       ctypes.c_float, fp, fp, fp, fp, ctypes.c_int]
        _LIB = L
    return _LIB


def _f(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_float))


class TightModel:
    """Takes the model as oracle.ref_numpy.load_model returns it (float32)."""

    def __init__(self, m):
        L = lib()
        c = lambda a: np.ascontiguousarray(a, dtype=np.float32)
        self.nk = int(m.joint_regressor.shape[1])
        par = np.ascontiguousarray(np.asarray(m.parents, dtype=np.int32))
        self._keep = [c(m.v_template), c(m.shapedirs), c(m.posedirs), c(m.J_regressor), c(m.weights), c(m.joint_regressor)]
        self.h = L.smpl_omp_create(*[_f(a) for a in self._keep], self.nk, par.ctypes.data_as(ctypes.POINTER(ctypes.c_int32)))
        if not self.h:
            raise ValueError('smpl_omp_create failed (more than 4 skinning weights on a vertex?)')

    def __del__(self):
        if getattr(self, 'h', None):
            lib().smpl_omp_free(self.h)
            self.h = None

    def forward(self, beta, theta, cam_pos, cam_rot, focal, ss, img_w, img_h, want_verts=True, nthreads=0, out=None):
        B = beta.shape[0]
        c = lambda a: np.ascontiguousarray(a, dtype=np.float32)
        beta, theta, cam_pos, cam_rot, focal = c(beta), c(theta), c(cam_pos), c(cam_rot), c(np.reshape(focal, [-1]))
        o = out or dict(verts=np.empty((B, 6890, 3), np.float32) if want_verts else None,
                        J_transformed=np.empty((B, 24, 3), np.float32), joints=np.empty((B, self.nk, 3), np.float32),
                        px=np.empty((B, 24, 2), np.float32))
        used = lib().smpl_omp_forward(self.h, B, _f(beta), _f(theta), _f(cam_pos), _f(cam_rot), _f(focal), ss, img_w, img_h,
                                      _f(o['verts']) if o['verts'] is not None else None, _f(o['J_transformed']),
                                      _f(o['joints']), _f(o['px']), nthreads)
        o['threads'] = used
        return o
