"""CPU: the C-ABI library loads and exports every symbol include/smplb200.h declares; the ctypes
table covers the header; argument validation that needs no GPU."""
import ctypes as C
import os
import re

from conftest import ROOT
from smplb200 import _lib


def header_symbols():
    src = open(os.path.join(ROOT, 'include', 'smplb200.h')).read()
    src = re.sub(r'/\*.*?\*/', '', src, flags=re.S)
    return sorted(set(re.findall(r'\b(smplb_[a-z0-9_]+)\s*\(', src)))


def test_library_exports_every_declared_symbol():
    syms = header_symbols()
    assert len(syms) >= 15
    l = C.CDLL(_lib.LIB_PATH)
    for s in syms:
        assert hasattr(l, s), 'libsmplb200.so does not export %s' % s


def test_ctypes_table_matches_header():
    assert sorted(_lib.SYMBOLS) == header_symbols()


def test_struct_layouts():
    assert C.sizeof(_lib.CamCfg) == 16
    assert C.sizeof(_lib.ForwardIO) == 8 + 12 * 8 + 16
    assert C.sizeof(_lib.BackwardIO) == 8 + 9 * 8 + 16 + 2 * 8


def test_argument_validation_without_gpu():
    L = _lib.lib()
    assert L.smplb_version().startswith(b'smplb200')
    h = C.c_void_p()
    assert L.smplb_model_create(C.byref(h), 0, None, None, None, None, None, None, 19, None) == _lib.E_INVALID_ARG
    assert b'null' in L.smplb_last_error()
    assert L.smplb_forward(None, None, None, 0, 0, None) == _lib.E_INVALID_ARG
    assert L.smplb_project(-1, 1, None, None, None, None, 1, None, None) == _lib.E_BAD_SHAPE
    assert L.smplb_adam_step(4, None, None, None, None, 0, 1e-3, 0.9, 0.999, 1e-8, None) == _lib.E_INVALID_ARG
    assert L.smplb_model_destroy(None) == 0
    # workspace size is pure arithmetic
    n1 = L.smplb_workspace_bytes(None, 1024, 0)
    n2 = L.smplb_workspace_bytes(None, 2048, 0)
    assert n1 >= 1024 * 20736 * 4 and n2 > n1
    assert L.smplb_workspace_bytes(None, 1024, _lib.F_NO_SKIN) < 1024 * 4096


def test_model_outside_fp16_range_is_rejected_before_any_cuda_call():
    """The tensor-core kernels split operands into an fp16 main part + bf16 residuals (csrc/common.cuh), so
    smplb_model_create refuses model constants that are not finite or exceed the fp16 range."""
    import numpy as np
    L = _lib.lib()
    f = lambda a: a.ctypes.data_as(C.POINTER(C.c_float))
    vt = np.zeros(6890 * 3, np.float32)
    sd = np.zeros(10 * 20670, np.float32)
    pd = np.zeros(207 * 20670, np.float32)
    jr = np.zeros(6890 * 24, np.float32)
    w = np.zeros(6890 * 24, np.float32)
    kr = np.zeros(6890 * 19, np.float32)
    par = np.array([-1, 0, 0, 0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 9, 9, 12, 13, 14, 16, 17, 18, 19, 20, 21], np.int32)
    for bad in (1e5, np.nan, -np.inf):
        pd[12345] = bad
        h = C.c_void_p()
        rc = L.smplb_model_create(C.byref(h), 0, f(vt), f(sd), f(pd), f(jr), f(w), f(kr), 19,
                                  par.ctypes.data_as(C.POINTER(C.c_int32)))
        assert rc == _lib.E_UNSUPPORTED and b'65504' in L.smplb_last_error()
