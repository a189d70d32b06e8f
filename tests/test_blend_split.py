"""Host-side check of the operand split the blend GEMM uses (csrc/blend_tc.cu, common.cuh):
    X.D ~= 2^-e [ f16(X).f16(D') + f16(X - f16(X)).f16(D') + f16(X).f16(D' - f16(D')) ],  D' = 2^e D with max|D'| in [2^14, 2^15)
(and the first version of this round, f16 main product + bf16 corrections on the unscaled D, for comparison)
emulated in NumPy with the same rounding rules (fp16 and bf16 round to nearest even, fp16 subnormals included;
the TF32 helpers model cvt.rna.tf32 for the 1x / 3xTF32 comparison rows) against float64, on operands shaped like [beta | R - I] x [shapedirs; posedirs]
(batch_smpl.py:91-95, :107-115).  The bound asserted here is what DESIGN.md quotes."""
import numpy as np


def tf32_rna(x):
    u = x.astype(np.float32).view(np.uint32).astype(np.uint64)
    return ((u + 0x1000) & 0xFFFFE000).astype(np.uint32).view(np.float32)


def bf16_rne(x):
    u = x.astype(np.float32).view(np.uint32).astype(np.uint64)
    return ((u + 0x7FFF + ((u >> 16) & 1)) & 0xFFFF0000).astype(np.uint32).view(np.float32)


def test_split_error_bounds():
    rng = np.random.default_rng(0)
    B, N = 256, 2048
    X = np.concatenate([np.clip(rng.standard_normal((B, 10)), -3, 3), rng.uniform(-2, 2, (B, 207))], 1).astype(np.float32)
    D = np.concatenate([rng.normal(0, 0.01, (10, N)), rng.normal(0, 0.005, (207, N))]).astype(np.float32)
    X[:, 50:60] *= 1e-5     # entries in the fp16 subnormal range: the residual part must carry them
    D[100:110] *= 1e-4
    f = lambda a: a.astype(np.float64)
    exact = f(X) @ f(D)
    Xh, Dh = tf32_rna(X), tf32_rna(D)
    Xl, Dl = X - Xh, D - Dh
    one = f(Xh) @ f(Dh)
    three = one + f(tf32_rna(Xl)) @ f(Dh) + f(Xh) @ f(tf32_rna(Dl))
    Xf, Df = X.astype(np.float16).astype(np.float32), D.astype(np.float16).astype(np.float32)
    mixed = (f(Xf) @ f(Df) + f((X - Xf).astype(np.float16).astype(np.float32)) @ f(Df) +
             f(bf16_rne(X)) @ f(bf16_rne(D - Df)))
    # the kernel's scheme: D scaled by a power of two so that both of its fp16 parts are normal numbers
    e2 = 15 - int(np.frexp(np.abs(D).max())[1])
    Dp = D * np.float32(2.0 ** e2)
    assert 2.0 ** 14 <= np.abs(Dp).max() < 2.0 ** 15
    Dph = Dp.astype(np.float16).astype(np.float32)
    Dpl = (Dp - Dph).astype(np.float16).astype(np.float32)
    Xl16 = (X - Xf).astype(np.float16).astype(np.float32)
    scaled = (f(Xf) @ f(Dph) + f(Xl16) @ f(Dph) + f(Xf) @ f(Dpl)) * 2.0 ** -e2
    scale = np.abs(exact).max()
    es = np.abs(scaled - exact).max() / scale
    assert es < 2e-6            # all-fp16 split with the scaled basis (what blend_tc.cu computes)
    e1, e3, em = (np.abs(r - exact).max() / scale for r in (one, three, mixed))
    assert es <= em * 1.05      # at least as good as the bf16-correction version it replaced
    assert e1 > 1e-4            # a single TF32 product is not enough (SURVEY.md header table)
    assert e3 < 5e-7            # 3xTF32
    assert em < 4e-6            # f16 main product + two bf16 corrections: relative to the OFFSET scale ...
    # ... and the offsets are an order of magnitude below the vertex scale the 1e-5 tolerance refers to
    assert em * scale / 1.0 < 1e-6


def test_split_parts_are_exact_in_their_formats():
    rng = np.random.default_rng(1)
    x = rng.standard_normal(4096).astype(np.float32)
    hi = tf32_rna(x)
    assert np.all((hi.view(np.uint32) & 0x1FFF) == 0)
    assert np.all(np.abs(x - hi) <= np.abs(x) * 2.0 ** -11)
    lo = bf16_rne(x - hi)
    assert np.all((lo.view(np.uint32) & 0xFFFF) == 0)
    assert np.all(np.abs((x - hi) - lo) <= np.abs(x) * 2.0 ** -19)
    h = x.astype(np.float16).astype(np.float32)          # the kernel's main part: same 11 significant bits
    big = np.abs(x) >= 2.0 ** -14
    assert np.all(np.abs(x - h)[big] <= np.abs(x)[big] * 2.0 ** -11)
    assert np.all(np.abs((x - h) - bf16_rne(x - h)) <= np.maximum(np.abs(x), 2.0 ** -14) * 2.0 ** -19)
