"""CPU tests of the oracle itself: known answers, analytic identities, golden fixtures, and the
independent torch twin.  (The reference holds no test for this path — SURVEY.md §4 — so these are
derived from its source.)"""
import os

import numpy as np
import pytest
import torch

from conftest import maxrel, ROOT
from oracle import ref_numpy as rn, ref_torch as rt
from smplb200 import synthetic as syn

GOLD = os.path.join(ROOT, 'tests', 'golden')


def test_project_known_answer():
    """project.py:92-97 example; expected values derived by hand in SURVEY.md §8c."""
    g = np.load(os.path.join(GOLD, 'project_kat.npz'))
    uv = rn.project(g['pts'], g['cam_pos'], g['cam_angles'], g['focal'])
    expect = np.array([[-0.05773503, 0.0], [-0.04013873, 0.01523884]])
    assert np.abs(uv - expect).max() < 5e-9
    assert np.abs(uv - g['uv']).max() == 0.0
    uv32 = rn.project(*(g[k].astype(np.float32) for k in ('pts', 'cam_pos', 'cam_angles', 'focal')))
    assert np.abs(uv32 - expect).max() < 1e-7


def test_rest_pose_identities(oracle64):
    m = oracle64
    o = rn.smpl_forward(m, np.zeros((2, 10)), np.zeros((2, 72)))
    wsum = m.weights.sum(axis=1, keepdims=True)                  # fp32-rounded weights: 1 +- 1e-7
    assert np.abs(o['verts'] - wsum * m.v_template).max() < 1e-12
    assert np.abs(o['Rs'] - np.eye(3)).max() < 1e-12
    J0 = m.J_regressor.T @ m.v_template
    assert np.abs(o['J_transformed'] - J0).max() < 1e-12
    assert np.abs(o['verts'] - m.v_template).max() < 1e-6


def test_root_rotation_only(oracle64):
    m = oracle64
    theta = np.zeros((1, 72))
    theta[0, :3] = [0.3, -1.1, 0.7]
    o = rn.smpl_forward(m, np.zeros((1, 10)), theta)
    R0 = o['Rs'][0, 0]
    J0 = (m.J_regressor.T @ m.v_template)[0]
    wsum = m.weights.sum(axis=1, keepdims=True)
    expect = wsum * ((m.v_template - J0) @ R0.T + J0)   # pose blend shapes ignore the root (batch_smpl.py:109)
    assert np.abs(o['verts'][0] - expect).max() < 1e-6
    assert np.abs(R0 @ R0.T - np.eye(3)).max() < 1e-7   # eps inside the norm (batch_lbs.py:45)


def test_two_rodrigues_agree():
    r = np.random.default_rng(0).normal(0, 0.8, size=(64, 3))
    assert np.abs(rn.batch_rodrigues(r) - rn.rodrigues_batch(r)).max() < 1e-7
    z = np.zeros((1, 3))
    assert np.abs(rn.rodrigues_batch(z) - np.eye(3)).max() == 0.0
    assert np.abs(rn.batch_rodrigues(z) - np.eye(3)).max() < 1e-12


def test_skew_layout():
    v = np.array([[1.0, 2.0, 3.0]])
    assert np.array_equal(rn.batch_skew(v)[0], np.array([[0, -3, 2], [3, 0, -1], [-2, 1, 0]], dtype=float))


def test_tileW_equals_contracted_and_fp32_floor(oracle64, oracle32):
    inp = syn.make_inputs(6, seed=3, dtype=np.float64)
    a = rn.smpl_forward(oracle64, inp['beta'], inp['theta'], tile_W=True)
    b = rn.smpl_forward(oracle64, inp['beta'], inp['theta'], tile_W=False)
    c = rn.smpl_forward(oracle32, inp['beta'], inp['theta'], tile_W=True)
    for k in ('verts', 'joints', 'J_transformed'):
        assert maxrel(b[k], a[k]) < 1e-14
        assert maxrel(c[k], a[k]) < 2e-6      # fp32 restatement floor (~2e-7, SURVEY header table)


@pytest.mark.parametrize('name', ['smpl_synth_seed0_B8.npz', 'smpl_surreal16.npz'])
def test_golden_fixtures(oracle64, name):
    g = np.load(os.path.join(GOLD, name))
    o = rn.smpl_forward(oracle64, g['beta'], g['theta'], tile_W=False)
    uv, px = rn.project_joints_px(o['J_transformed'], g['cam_pos'], g['cam_rot'], g['focal'])
    assert maxrel(o['verts'][:, g['vsub']], g['verts_sub']) < 1e-13
    assert maxrel(o['verts'].sum(axis=1), g['verts_sum']) < 1e-12
    assert maxrel(o['joints'], g['joints']) < 1e-13
    assert maxrel(o['J_transformed'], g['J_transformed']) < 1e-13
    assert maxrel(o['A'][:, :, :3, :], g['A']) < 1e-13
    assert maxrel(px, g['px']) < 1e-12


def test_torch_twin_matches_numpy(synth_dd, oracle64):
    t64 = rt.load_model(synth_dd, dtype=torch.float64)
    inp = syn.make_inputs(5, seed=11, dtype=np.float64)
    a = rn.smpl_forward(oracle64, inp['beta'], inp['theta'])
    b = rt.smpl_forward(t64, torch.from_numpy(inp['beta']), torch.from_numpy(inp['theta']))
    for k in ('verts', 'joints', 'J_transformed', 'A'):
        assert maxrel(b[k].numpy(), a[k]) < 1e-13
    _, px_n = rn.project_joints_px(a['J_transformed'], inp['cam_pos'], inp['cam_rot'], inp['focal'])
    _, px_t = rt.project_joints_px(b['J_transformed'], *(torch.from_numpy(inp[k]) for k in ('cam_pos', 'cam_rot', 'focal')))
    assert maxrel(px_t.numpy(), px_n) < 1e-13
    ln, _, _ = rn.reproj_loss(a['J_transformed'], a['J_transformed'] * 1.01, inp['cam_pos'], inp['cam_rot'], inp['focal'])
    lt, _, _ = rt.reproj_loss(b['J_transformed'], b['J_transformed'] * 1.01, *(torch.from_numpy(inp[k]) for k in ('cam_pos', 'cam_rot', 'focal')))
    assert abs(float(lt) - float(ln)) < 1e-12 * max(1.0, abs(float(ln)))


def test_autograd_vs_finite_difference(synth_dd):
    """Gradient of the (uncapped) re-projection + 3D losses wrt theta/beta: autograd == central FD."""
    t64 = rt.load_model(synth_dd, dtype=torch.float64)
    inp = syn.make_inputs(2, seed=5, dtype=np.float64)
    gt = syn.make_inputs(2, seed=6, dtype=np.float64)
    cam = [torch.from_numpy(inp[k]) for k in ('cam_pos', 'cam_rot', 'focal')]
    beta = torch.from_numpy(inp['beta']).requires_grad_(True)
    theta = torch.from_numpy(inp['theta']).requires_grad_(True)
    with torch.no_grad():
        g = rt.smpl_forward(t64, torch.from_numpy(inp['beta']), torch.from_numpy(gt['theta']), tile_W=False)

    def loss_fn(b, t):
        o = rt.smpl_forward(t64, b, t, tile_W=False)
        l2d, _, _ = rt.reproj_loss(o['J_transformed'], g['J_transformed'], *cam, cap=False)
        return l2d + rt.loss3d(g['verts'], o['verts'], g['J_transformed'], o['J_transformed'])

    L = loss_fn(beta, theta)
    L.backward()
    rng = np.random.default_rng(0)
    for _ in range(6):
        i, j = rng.integers(0, 2), rng.integers(0, 72)
        h = 1e-6
        tp = theta.detach().clone(); tp[i, j] += h
        tm = theta.detach().clone(); tm[i, j] -= h
        fd = (loss_fn(beta.detach(), tp) - loss_fn(beta.detach(), tm)) / (2 * h)
        assert abs(float(fd) - float(theta.grad[i, j])) < 1e-6 * max(1.0, abs(float(fd)))
    for _ in range(3):
        i, j = rng.integers(0, 2), rng.integers(0, 10)
        h = 1e-6
        bp = beta.detach().clone(); bp[i, j] += h
        bm = beta.detach().clone(); bm[i, j] -= h
        fd = (loss_fn(bp, theta.detach()) - loss_fn(bm, theta.detach())) / (2 * h)
        assert abs(float(fd) - float(beta.grad[i, j])) < 1e-6 * max(1.0, abs(float(fd)))


def test_adam_tf1_hand_trace():
    x = np.array([1.0, -2.0]); g = np.array([0.5, -0.25])
    m = np.zeros(2); v = np.zeros(2)
    x1, m1, v1 = rn.adam_tf1(x, g, m, v, 1, lr=0.1)
    # t=1: m=0.1g, v=0.001g^2, lr_t = 0.1*sqrt(0.001)/0.1 ; step = lr_t*m/(sqrt(v)+eps) ~= 0.1*sign(g)
    assert np.allclose(x1, x - 0.1 * np.sign(g), atol=1e-6)
    x2, m2, v2 = rn.adam_tf1(x1, g, m1, v1, 2, lr=0.1)
    lr2 = 0.1 * np.sqrt(1 - 0.999 ** 2) / (1 - 0.9 ** 2)
    assert np.allclose(x2, x1 - lr2 * m2 / (np.sqrt(v2) + 1e-8))


def test_orth_proj_and_pixels():
    X = np.arange(24, dtype=np.float64).reshape(2, 4, 3)
    cam = np.array([[2.0, 1.0, -1.0], [0.5, 0.0, 3.0]])
    out = rn.batch_orth_proj_idrot(X, cam)
    assert np.allclose(out[0, 1], 2.0 * (X[0, 1, :2] + [1.0, -1.0]))
    px = rn.to_pixels(np.array([[0.0, 0.0], [0.012, -0.012]]))
    assert np.allclose(px, [[160.0, 120.0], [160.0 + 120.0, 120.0 - 120.0]])


def test_rotate_global_pose_composes_rotations():
    """utils.py:152-199: result rotation == Ry(zrot+pi/2) Rx(-pi/2) R(theta_0)."""
    rng = np.random.default_rng(2)
    th = rng.normal(0, 0.5, size=(4, 72)); zrot = rng.uniform(-1, 1, size=4)
    out = rn.rotate_global_pose(th, zrot)
    R_out = rn.rodrigues_batch(out[:, :3])
    Rx = rn.rodrigues_batch(np.tile([[-np.pi / 2, 0, 0]], (4, 1)))
    Ry = rn.rodrigues_batch(np.stack([np.zeros(4), zrot + np.pi / 2, np.zeros(4)], 1))
    R0 = rn.rodrigues_batch(th[:, :3])
    assert np.abs(R_out - Ry @ Rx @ R0).max() < 1e-9
    assert np.array_equal(out[:, 3:], th[:, 3:])
