"""smplb_cam_losses_static: the allocation-free variant of smplb_cam_losses (scratch in a fixed device array inside the
library, no counterpart in the reference) against the existing entry point, through the C ABI, on two calls in a row
(the array keeps the previous call's contents).  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _inputs(B, seed, dev):
    rng = np.random.default_rng(seed)
    f = lambda a: torch.tensor(np.asarray(a, dtype=np.float32), device=dev)
    j3 = f(rng.normal(0, 0.4, size=(B, 24, 3)))
    j2 = f(rng.uniform(20, 220, size=(B, 24, 2)))
    cp = f(np.array([0.0, -0.35, -4.0]) + rng.normal(0, 0.1, size=(B, 3)))
    cr = f(rng.normal(0, 0.05, size=(B, 3)))
    cf = f(np.full(B, 0.05) + rng.normal(0, 0.002, size=B))
    fr = f(rng.normal(0.05, 0.01, size=B))
    return j3, j2, cp, cr, cf, fr


def _call(L, static, B, inp, dev):
    from smplb200._lib import CamCfg, check
    j3, j2, cp, cr, cf, fr = inp
    p = lambda t: C.c_void_p(t.data_ptr())
    out = {k: torch.full(s, 7.0, device=dev) for k, s in (('loss', (4,)), ('g_cp', (B, 3)), ('g_cr', (B, 3)), ('g_cf', (B,)), ('g_fr', (B,)))}
    cfg = CamCfg(320.0, 240.0, 0.024, 0.05)
    st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    args = [B, p(j3), p(j2), p(cp), p(cr), p(cf), p(fr), C.byref(cfg), 10.0, p(out['loss']), p(out['g_cp']), p(out['g_cr']),
            p(out['g_cf']), p(out['g_fr'])]
    if static:
        check(L.smplb_cam_losses_static(*args, st))
    else:
        scratch = torch.empty(max(int(L.smplb_cam_losses_scratch_bytes(B)), 16), dtype=torch.uint8, device=dev)
        check(L.smplb_cam_losses(*args, p(scratch), st))
    torch.cuda.synchronize()
    return {k: v.cpu().numpy() for k, v in out.items()}


def test_cam_losses_static_matches_the_scratch_version_twice():
    from smplb200 import _lib
    L = _lib.lib()
    dev = torch.device('cuda', 0)
    with torch.cuda.device(dev):
        # two calls in a row with different sizes: the second (smaller) call must not see the first one's partial sums
        for B, seed in ((1000, 1), (37, 2), (1000, 3)):
            inp = _inputs(B, seed, dev)
            want = _call(L, False, B, inp, dev)
            got = _call(L, True, B, inp, dev)
            for k in want:
                scale = max(np.abs(want[k]).max(), 1e-30)
                assert np.abs(got[k] - want[k]).max() / scale <= 1e-5, k      # the tolerance of the other GPU tests
            assert np.isfinite(got['loss']).all() and got['loss'][3] > 0


def test_cam_losses_static_rejects_sizes_beyond_the_array():
    from smplb200 import _lib
    L = _lib.lib()
    dev = torch.device('cuda', 0)
    cfg = _lib.CamCfg(320.0, 240.0, 0.024, 0.05)
    loss = torch.zeros(4, device=dev)
    rc = L.smplb_cam_losses_static(65537, None, None, None, None, None, None, C.byref(cfg), 10.0,
                                   C.c_void_p(loss.data_ptr()), None, None, None, None, None)
    assert rc == _lib.E_UNSUPPORTED
    assert b'SMPLB_CAM_LOSSES_STATIC_MAX_B' in L.smplb_last_error()
