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
