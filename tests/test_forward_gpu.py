"""GPU parity: the CUDA path (through the C ABI) against the float64 oracle on the same seeded inputs.
Tolerances (BASELINE.json north_star / SURVEY.md §8c): max|gpu-oracle| / max|oracle| <= 1e-5 for
vertices, joints, keypoints and 2D keypoints."""
import os

import numpy as np
import pytest
import torch

from conftest import maxrel, ROOT
from oracle import ref_numpy as rn
from smplb200 import synthetic as syn, _lib
import smplb200

pytestmark = pytest.mark.gpu
TOL = 1e-5
GOLD = os.path.join(ROOT, 'tests', 'golden')
# CUDA-core verification kernels, tcgen05 kernels (f16 main product + f16/bf16 corrections), fused single-kernel forward
BLENDS = [_lib.F_BLEND_SIMT | _lib.F_SKIN_SIMT, 0, _lib.F_FUSED]


def _check(out, o, px=None, tol=TOL):
    assert maxrel(out['verts'].cpu().numpy(), o['verts']) <= tol
    assert maxrel(out['joints'].cpu().numpy(), o['J_transformed']) <= tol
    assert maxrel(out['kp'].cpu().numpy(), o['joints']) <= tol
    assert maxrel(out['Rs'].cpu().numpy(), o['Rs']) <= tol
    if px is not None:
        assert maxrel(out['px'].cpu().numpy(), px) <= tol


@pytest.mark.parametrize('flags', BLENDS)
@pytest.mark.parametrize('B', [1, 3, 37, 64, 300])
def test_forward_matches_oracle(gpu_model, oracle64, B, flags):
    inp = syn.make_inputs(B, seed=1000 + B, dtype=np.float32)
    out = gpu_model.forward(inp['beta'], inp['theta'], inp['cam_pos'], inp['cam_rot'], inp['focal'],
                            want_A=True, flags=flags)
    f64 = {k: v.astype(np.float64) for k, v in inp.items()}
    o = rn.smpl_forward(oracle64, f64['beta'], f64['theta'], tile_W=False)
    uv, px = rn.project_joints_px(o['J_transformed'], f64['cam_pos'], f64['cam_rot'], f64['focal'])
    _check(out, o, px)
    assert maxrel(out['uv'].cpu().numpy(), uv) <= TOL
    assert maxrel(out['A'].cpu().numpy(), o['A'][:, :, :3, :]) <= TOL


@pytest.mark.parametrize('flags', BLENDS)
@pytest.mark.parametrize('name', ['smpl_synth_seed0_B8.npz', 'smpl_surreal16.npz'])
def test_forward_matches_golden(gpu_model, name, flags):
    g = np.load(os.path.join(GOLD, name))
    out = gpu_model.forward(g['beta'].astype(np.float32), g['theta'].astype(np.float32),
                            g['cam_pos'].astype(np.float32), g['cam_rot'].astype(np.float32),
                            g['focal'].astype(np.float32), want_A=True, flags=flags)
    v = out['verts'].cpu().numpy()
    assert maxrel(v[:, g['vsub']], g['verts_sub']) <= TOL
    assert maxrel(v.astype(np.float64).sum(axis=1), g['verts_sum']) <= 1e-4 * 10  # sum of 6890 terms
    assert maxrel(out['joints'].cpu().numpy(), g['J_transformed']) <= TOL
    assert maxrel(out['kp'].cpu().numpy(), g['joints']) <= TOL
    assert maxrel(out['Rs'].cpu().numpy(), g['Rs']) <= TOL
    assert maxrel(out['A'].cpu().numpy(), g['A']) <= TOL
    assert maxrel(out['px'].cpu().numpy(), g['px']) <= TOL


@pytest.mark.parametrize('flags', BLENDS)
def test_rest_pose_and_zero_theta(gpu_model, oracle64, flags):
    """theta = 0 exercises the eps-in-norm Rodrigues (batch_lbs.py:45-46): finite, R = I."""
    out = gpu_model.forward(np.zeros((2, 10), np.float32), np.zeros((2, 72), np.float32), flags=flags)
    m = oracle64
    wsum = m.weights.sum(axis=1, keepdims=True)
    assert torch.isfinite(out['verts']).all()
    assert maxrel(out['verts'].cpu().numpy()[0], wsum * m.v_template) <= TOL
    assert np.abs(out['Rs'].cpu().numpy() - np.eye(3)).max() <= 1e-7


def test_dense_lbs_weights(oracle64):
    """All 24 weights non-zero per vertex: handled by the tensor-core skinning (dense W x A matmul,
    batch_smpl.py:125-127); only the compressed-weights verification path refuses it."""
    dd = syn.make_synthetic_smpl(3, dense_weights=True)
    model = smplb200.Model(dd, device=0)
    inp = syn.make_inputs(45, seed=11, dtype=np.float32)
    out = model.forward(inp['beta'], inp['theta'])
    m64 = rn.load_model(dd, dtype=np.float64)
    o = rn.smpl_forward(m64, inp['beta'].astype(np.float64), inp['theta'].astype(np.float64), tile_W=False)
    _check(out, o)
    with pytest.raises(smplb200.SmplbError) as ei:
        model.forward(inp['beta'], inp['theta'], flags=_lib.F_SKIN_SIMT)
    assert ei.value.code == _lib.E_UNSUPPORTED


def test_empty_batch_and_errors(gpu_model):
    out = gpu_model.forward(np.zeros((0, 10), np.float32), np.zeros((0, 72), np.float32))
    assert out['verts'].shape == (0, 6890, 3)
    with pytest.raises(ValueError):
        gpu_model.forward(np.zeros((2, 9), np.float32), np.zeros((2, 72), np.float32))
    with pytest.raises(smplb200.SmplbError) as ei:   # verts requested together with NO_SKIN
        _raise_no_skin(gpu_model)
    assert ei.value.code == _lib.E_INVALID_ARG


def _raise_no_skin(model):
    import ctypes as C
    io = _lib.ForwardIO()
    io.B = 1
    t = torch.zeros(100, device='cuda')
    io.beta = t.data_ptr(); io.theta = t.data_ptr(); io.verts = t.data_ptr()
    ws = torch.empty(1 << 20, dtype=torch.uint8, device='cuda')
    _lib.check(_lib.lib().smplb_forward(model._h, C.byref(io), C.c_void_p(ws.data_ptr()), ws.numel(),
                                        _lib.F_NO_SKIN, None))


def test_joints_only_path(gpu_model, oracle64):
    inp = syn.make_inputs(50, seed=77, dtype=np.float32)
    out = gpu_model.forward(inp['beta'], inp['theta'], inp['cam_pos'], inp['cam_rot'], None,
                            flags=_lib.F_NO_SKIN)
    o = rn.smpl_forward(oracle64, inp['beta'].astype(np.float64), inp['theta'].astype(np.float64), tile_W=False)
    _, px = rn.project_joints_px(o['J_transformed'], inp['cam_pos'].astype(np.float64),
                                 inp['cam_rot'].astype(np.float64), np.full(50, 0.05))
    assert 'verts' not in out
    assert maxrel(out['joints'].cpu().numpy(), o['J_transformed']) <= TOL
    assert maxrel(out['px'].cpu().numpy(), px) <= TOL


def test_large_pose_angles(gpu_model, oracle64):
    """SURREAL-like extremes: |theta_j| up to ~4 rad."""
    rng = np.random.default_rng(9)
    theta = rng.normal(0, 1.5, size=(32, 72)).astype(np.float32)
    beta = rng.uniform(-3, 3, size=(32, 10)).astype(np.float32)
    out = gpu_model.forward(beta, theta)
    o = rn.smpl_forward(oracle64, beta.astype(np.float64), theta.astype(np.float64), tile_W=False)
    _check(out, o)


def test_fullsize_properties(gpu_model):
    """B = 16384 (too big for the oracle): size-independent properties.
    (i) samples are independent: a sub-batch reproduces the same rows bit for bit;
    (ii) root-rotation equivariance: rotating theta_0 by Q rotates posed joints about J_0 ... checked
         through orthonormal Rs and |verts| bounded; (iii) keypoints == regressor applied to verts."""
    B = 16384
    inp = syn.make_inputs(B, seed=4242, dtype=np.float32)
    out = gpu_model.forward(inp['beta'], inp['theta'], inp['cam_pos'], inp['cam_rot'], inp['focal'])
    verts = out['verts'].clone(); kp = out['kp'].clone(); px = out['px'].clone()
    assert torch.isfinite(verts).all() and torch.isfinite(px).all()
    idx = torch.arange(5000, 5000 + 97)
    sub = gpu_model.forward(inp['beta'][idx.numpy()], inp['theta'][idx.numpy()], inp['cam_pos'][idx.numpy()],
                            inp['cam_rot'][idx.numpy()], inp['focal'][idx.numpy()])
    assert torch.equal(sub['verts'], verts[5000:5097])
    assert torch.equal(sub['px'], px[5000:5097])
    Rs = out['Rs']
    eye = torch.eye(3, device=Rs.device)
    assert (Rs @ Rs.transpose(-1, -2) - eye).abs().max() < 1e-5
    jr = torch.from_numpy(gpu_model.joint_regressor).to(verts.device).double()
    kp_ref = torch.einsum('bvc,vk->bkc', verts[:256].double(), jr)
    assert (kp[:256].double() - kp_ref).abs().max() / kp_ref.abs().max() < 1e-6


def test_standalone_ops(gpu_model, oracle64):
    rng = np.random.default_rng(5)
    r = rng.normal(0, 0.9, size=(1000, 3)).astype(np.float32)
    r[0] = 0.0
    R0 = smplb200.rodrigues(r, 0).cpu().numpy()
    R1 = smplb200.rodrigues(r, 1).cpu().numpy()
    assert maxrel(R0, rn.batch_rodrigues(r.astype(np.float64))) <= TOL
    assert maxrel(R1, rn.rodrigues_batch(r.astype(np.float64))) <= TOL
    assert np.array_equal(R1[0], np.eye(3, dtype=np.float32))
    # project: reference-style tiled cameras, KAT first
    g = np.load(os.path.join(GOLD, 'project_kat.npz'))
    uv = smplb200.project(g['pts'], g['cam_pos'], g['cam_angles'], g['focal']).cpu().numpy()
    assert np.abs(uv - np.array([[-0.05773503, 0.0], [-0.04013873, 0.01523884]])).max() < 2e-8
    pts = rng.normal(0, 1, size=(480, 3)).astype(np.float32)
    cp = np.tile(np.array([[0.1, -0.3, -4.0]], np.float32), (480, 1)) + rng.normal(0, 0.05, size=(480, 3)).astype(np.float32)
    ca = rng.normal(0, 0.1, size=(480, 3)).astype(np.float32)
    fl = np.full(480, 0.05, np.float32)
    uv = smplb200.project(pts, cp, ca, fl).cpu().numpy()
    assert maxrel(uv, rn.project(*(x.astype(np.float64) for x in (pts, cp, ca, fl)))) <= TOL
    uv2 = smplb200.project(pts, cp[::24].copy(), ca[::24].copy(), fl[::24].copy(), P=24, cam_tiled=False).cpu().numpy()
    ref2 = rn.project(pts.astype(np.float64), rn.tile_cam(cp[::24].astype(np.float64)), rn.tile_cam(ca[::24].astype(np.float64)),
                      rn.tile_cam(fl[::24].astype(np.float64)))
    assert maxrel(uv2, ref2) <= TOL
    # global rigid transformation incl. rotate_base
    Rs = rn.batch_rodrigues(rng.normal(0, 0.5, size=(24 * 7, 3))).reshape(7, 24, 3, 3)
    Js = rng.normal(0, 0.4, size=(7, 24, 3))
    for rb in (False, True):
        nJ, A = smplb200.global_rigid(Rs.astype(np.float32), Js.astype(np.float32), oracle64.parents, rb)
        nJ_ref, A_ref = rn.batch_global_rigid_transformation(Rs, Js, oracle64.parents, rb)
        assert maxrel(nJ.cpu().numpy(), nJ_ref) <= TOL
        assert maxrel(A.cpu().numpy(), A_ref) <= TOL
    X = rng.normal(size=(5, 19, 3)).astype(np.float32); cam = rng.normal(size=(5, 3)).astype(np.float32)
    assert maxrel(smplb200.orth_proj(X, cam).cpu().numpy(), rn.batch_orth_proj_idrot(X.astype(np.float64), cam.astype(np.float64))) <= 2e-6
    # loss + adam
    a = rng.uniform(0, 320, size=(16, 24, 2)).astype(np.float32); b = a + rng.normal(0, 3, size=a.shape).astype(np.float32)
    l = smplb200.reproj_loss(a, b).cpu().numpy()
    ref = rn.mse(b.astype(np.float64), a.astype(np.float64), 0.024)
    assert abs(l[0] - ref) <= 1e-5 * ref and abs(l[1] - min(ref, 1 / 24)) <= 1e-5 * ref
    x = torch.tensor(rng.normal(size=100).astype(np.float32), device='cuda'); gg = torch.tensor(rng.normal(size=100).astype(np.float32), device='cuda')
    m = torch.zeros_like(x); v = torch.zeros_like(x)
    xn, mn, vn = x.cpu().numpy().astype(np.float64), np.zeros(100), np.zeros(100)
    for t in (1, 2, 3):
        smplb200.adam_step(x, gg, m, v, t)
        xn, mn, vn = rn.adam_tf1(xn, gg.cpu().numpy().astype(np.float64), mn, vn, t)
    assert maxrel(x.cpu().numpy(), xn) <= 2e-6


def test_dropin_smpl_class(synth_dd, oracle64, tmp_path):
    """Reads like the reference's own usage (tf_smpl/test_runner.py:19-29)."""
    from tf_smpl.batch_smpl import SMPL
    from tf_perspective_projection.project import project
    path = str(tmp_path / 'model_neutral_np.pkl')
    syn.write_pkl(path, synth_dd)
    smpl_model = SMPL(path)
    betas = (np.random.default_rng(0).standard_normal((4, 10)) * 0.1).astype(np.float32)
    thetas = (np.random.default_rng(1).standard_normal((4, 72)) * 0.2).astype(np.float32)
    verts, joints, Rs = smpl_model(betas, thetas, get_skin=True)
    o = rn.smpl_forward(oracle64, betas.astype(np.float64), thetas.astype(np.float64))
    assert maxrel(verts.cpu().numpy(), o['verts']) <= TOL
    assert maxrel(joints.cpu().numpy(), o['joints']) <= TOL
    assert maxrel(smpl_model.J_transformed.cpu().numpy(), o['J_transformed']) <= TOL
    only_joints = smpl_model(betas, thetas)
    assert torch.equal(only_joints, joints)
    lsp = SMPL(path, joint_type='lsp')
    assert lsp(betas, thetas).shape == (4, 14, 3)
    cam_p = np.tile([[0.0, -0.35, -4.0]], (96, 1)).astype(np.float32)
    cam_r = np.tile([[0.02, 0.05, -0.01]], (96, 1)).astype(np.float32)
    uv = project(smpl_model.J_transformed.reshape(-1, 3), cam_p, cam_r, np.full(96, 0.05, np.float32))
    ref = rn.project(o['J_transformed'].reshape(-1, 3), cam_p.astype(np.float64), cam_r.astype(np.float64), np.full(96, 0.05))
    assert maxrel(uv.cpu().numpy(), ref) <= TOL


def test_forward_host_chunked_pipeline_matches_device_call(gpu_model):
    """smplb_forward_host (host buffers, chunked H2D / kernels / D2H pipeline, 8192-sample chunks) returns exactly
    what the device-resident call computes, across a chunk boundary and with a ragged tail."""
    import ctypes as C
    import torch
    from smplb200 import _lib, synthetic as syn
    B = 8192 + 8192 + 777
    inp = syn.make_inputs(B, seed=77)
    out = gpu_model.forward(inp['beta'], inp['theta'], inp['cam_pos'], inp['cam_rot'], inp['focal'])
    torch.cuda.synchronize()
    NK = gpu_model.NK
    hb = {k: torch.from_numpy(v).pin_memory() for k, v in inp.items()}
    ho = {'verts': torch.empty((B, 6890, 3)).pin_memory(), 'joints': torch.empty((B, 24, 3)).pin_memory(),
          'kp': torch.empty((B, NK, 3)).pin_memory(), 'px': torch.empty((B, 24, 2)).pin_memory()}
    dbuf = {k: torch.empty_like(v, device='cuda') for k, v in {**hb, **ho}.items()}
    cfg = _lib.CamCfg(320.0, 240.0, 0.024, 0.05)
    FIO = _lib.ForwardIO
    hio = FIO(B, hb['beta'].data_ptr(), hb['theta'].data_ptr(), ho['verts'].data_ptr(), ho['joints'].data_ptr(),
              ho['kp'].data_ptr(), None, None, hb['cam_pos'].data_ptr(), hb['cam_rot'].data_ptr(),
              hb['focal'].data_ptr(), None, ho['px'].data_ptr(), cfg)
    dio = FIO(B, dbuf['beta'].data_ptr(), dbuf['theta'].data_ptr(), dbuf['verts'].data_ptr(), dbuf['joints'].data_ptr(),
              dbuf['kp'].data_ptr(), None, None, dbuf['cam_pos'].data_ptr(), dbuf['cam_rot'].data_ptr(),
              dbuf['focal'].data_ptr(), None, dbuf['px'].data_ptr(), cfg)
    ws, _ = gpu_model.workspace(B, 0)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    _lib.check(_lib.lib().smplb_forward_host(gpu_model._h, C.byref(hio), C.byref(dio), C.c_void_p(ws.data_ptr()),
                                             ws.numel(), 0, st))
    for k in ('verts', 'joints', 'kp', 'px'):
        assert torch.equal(ho[k], out[k].cpu()), k


def test_blend_split_precision(gpu_model, oracle64):
    """The blend GEMM evaluates X.D as f16(X).f16(D) + f16(Xl).f16(D) + bf16(X).bf16(Dl) (blend_tc.cu):
    worst-case inputs for the split (|beta| = 3 everywhere, large joint rotations so the pose feature R - I
    reaches its range [-2, 2]) must stay 5x inside the 1e-5 budget."""
    B = 384
    rng = np.random.default_rng(77)
    beta = (3.0 * rng.choice([-1.0, 1.0], size=(B, 10))).astype(np.float32)
    theta = rng.uniform(-2.5, 2.5, size=(B, 72)).astype(np.float32)
    out = gpu_model.forward(beta, theta, flags=0)
    o = rn.smpl_forward(oracle64, beta.astype(np.float64), theta.astype(np.float64), tile_W=False)
    assert maxrel(out['verts'].cpu().numpy(), o['verts']) <= 2e-6


def _oracle_chunked(oracle64, inp, chunk=256):
    """float64 oracle over a large batch in chunks (memory)."""
    outs = []
    for b0 in range(0, inp['beta'].shape[0], chunk):
        f64 = {k: v[b0:b0 + chunk].astype(np.float64) for k, v in inp.items()}
        o = rn.smpl_forward(oracle64, f64['beta'], f64['theta'], tile_W=False)
        _, px = rn.project_joints_px(o['J_transformed'], f64['cam_pos'], f64['cam_rot'], f64['focal'])
        outs.append((o['verts'], o['J_transformed'], o['joints'], px))
    return [np.concatenate(x) for x in zip(*outs)]


@pytest.mark.parametrize('flags', [0, _lib.F_FUSED])
def test_baseline_config2_B1024_matches_oracle(gpu_model, oracle64, flags):
    """BASELINE.json configs[1]: batched SMPL forward + perspective projection at exactly B = 1024, every sample against
    the float64 oracle (the largest oracle-checked batch was 300)."""
    B = 1024
    inp = syn.make_inputs(B, seed=4242, dtype=np.float32)
    out = gpu_model.forward(inp['beta'], inp['theta'], inp['cam_pos'], inp['cam_rot'], inp['focal'], flags=flags,
                            check_inputs=True)
    verts, joints, kp, px = _oracle_chunked(oracle64, inp)
    assert maxrel(out['verts'].cpu().numpy(), verts) <= TOL
    assert maxrel(out['joints'].cpu().numpy(), joints) <= TOL
    assert maxrel(out['kp'].cpu().numpy(), kp) <= TOL
    assert maxrel(out['px'].cpu().numpy(), px) <= TOL


def test_fused_forward_ragged_and_keypoints_only(gpu_model, oracle64):
    """SMPLB_F_FUSED: batches that end inside a 96-sample chunk / a 4-sample T group, and the get_skin=False use
    (keypoints without a verts buffer: the compact keypoint rows come from the fused epilogue)."""
    for B in (1, 5, 97, 191, 200):
        inp = syn.make_inputs(B, seed=77 + B, dtype=np.float32)
        f64 = {k: v.astype(np.float64) for k, v in inp.items()}
        o = rn.smpl_forward(oracle64, f64['beta'], f64['theta'], tile_W=False)
        out = gpu_model.forward(inp['beta'], inp['theta'], flags=_lib.F_FUSED, out={})
        assert maxrel(out['verts'].cpu().numpy(), o['verts']) <= TOL
        assert maxrel(out['kp'].cpu().numpy(), o['joints']) <= TOL
        out2 = gpu_model.forward(inp['beta'], inp['theta'], want_verts=False, flags=_lib.F_FUSED, out={})
        assert 'verts' not in out2 or out2.get('verts') is None
        assert maxrel(out2['kp'].cpu().numpy(), o['joints']) <= TOL
    with pytest.raises(smplb200.SmplbError):      # the backward needs v_posed: not with the fused forward
        gpu_model.forward(inp['beta'], inp['theta'], flags=_lib.F_FUSED | _lib.F_KEEP_VPOSED, out={})


def test_input_range_check(gpu_model):
    """|beta| beyond the fp16 range cannot be carried by the split-operand GEMMs: smplb_forward_status reports
    SMPLB_E_NOT_FINITE instead of silently returning inf; in-range inputs pass."""
    inp = syn.make_inputs(8, seed=3, dtype=np.float32)
    gpu_model.forward(inp['beta'], inp['theta'], check_inputs=True)
    bad = inp['beta'].copy()
    bad[5, 2] = 1.0e5
    with pytest.raises(smplb200.SmplbError) as e:
        gpu_model.forward(bad, inp['theta'], check_inputs=True)
    assert e.value.code == _lib.E_NOT_FINITE
    th = inp['theta'].copy()
    th[1, 7] = np.nan
    with pytest.raises(smplb200.SmplbError) as e:
        gpu_model.forward(inp['beta'], th, check_inputs=True)
    assert e.value.code == _lib.E_NOT_FINITE
    gpu_model.forward(inp['beta'], inp['theta'], check_inputs=True)      # the flag is cleared by the next call


def test_tolerance_report(gpu_model, oracle64, oracle32):
    """SURVEY 8c asks for three numbers side by side: the max-norm error (the acceptance criterion), the share of
    elements inside |a-b| <= 1e-5 |b| + 1e-5 max|b|, and the error of the float32 restatement of the reference itself
    (the floor).  Written to gpurun_out/tolerance_report.json (committed copy: profiles/r02_tolerance_report.json)."""
    import json
    B = 256
    inp = syn.make_inputs(B, seed=99, dtype=np.float32)
    f64 = {k: v.astype(np.float64) for k, v in inp.items()}
    o64 = rn.smpl_forward(oracle64, f64['beta'], f64['theta'], tile_W=False)
    _, px64 = rn.project_joints_px(o64['J_transformed'], f64['cam_pos'], f64['cam_rot'], f64['focal'])
    o32 = rn.smpl_forward(oracle32, inp['beta'], inp['theta'], tile_W=False)
    _, px32 = rn.project_joints_px(o32['J_transformed'], inp['cam_pos'], inp['cam_rot'], inp['focal'])
    rep = {}
    for name, fl in (('tcgen05 unfused', 0), ('tcgen05 fused', _lib.F_FUSED), ('CUDA-core verification', BLENDS[0])):
        out = gpu_model.forward(inp['beta'], inp['theta'], inp['cam_pos'], inp['cam_rot'], inp['focal'], flags=fl, out={})
        row = {}
        for key, got, want, floor in (('verts', out['verts'], o64['verts'], o32['verts']),
                                      ('J_transformed', out['joints'], o64['J_transformed'], o32['J_transformed']),
                                      ('keypoints', out['kp'], o64['joints'], o32['joints']),
                                      ('px', out['px'], px64, px32)):
            a = got.cpu().numpy().astype(np.float64)
            mixed = np.abs(a - want) <= 1e-5 * np.abs(want) + 1e-5 * np.abs(want).max()
            row[key] = {'max_norm': maxrel(a, want), 'inside_rtol_atol': float(mixed.mean()),
                        'fp32_restatement_max_norm': maxrel(floor, want)}
            assert row[key]['max_norm'] <= TOL and row[key]['inside_rtol_atol'] == 1.0
        rep[name] = row
    os.makedirs(os.path.join(ROOT, 'gpurun_out'), exist_ok=True)
    with open(os.path.join(ROOT, 'gpurun_out', 'tolerance_report.json'), 'w') as f:
        json.dump({'B': B, 'criterion': 'max|a-b|/max|b| <= 1e-5; also |a-b| <= 1e-5|b| + 1e-5 max|b| per element', 'paths': rep}, f, indent=1)
