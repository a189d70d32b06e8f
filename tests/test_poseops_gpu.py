"""GPU parity of the N2/N3 kernels (csrc/pose_ops.cu) through the C ABI against the float64 oracle:
values within 1e-5 (max-norm relative), gradients within 1e-4 (BASELINE.json north_star tolerances)."""
import numpy as np
import pytest

torch = pytest.importorskip('torch')
pytestmark = pytest.mark.gpu

from oracle import ref_numpy as rn, ref_torch as rt  # noqa: E402
from test_poseops_oracle import cam_inputs  # noqa: E402


def rel(a, b):
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


def test_add_axis_angle_and_zero_branches():
    from smplb200 import api
    rng = np.random.default_rng(0)
    a = rng.normal(0, 1.0, size=(257, 3))
    b = rng.normal(0, 1.0, size=(257, 3))
    a[3] = 0
    b[5] = 0
    a[7] = b[7] = 0
    got = api.add_axis_angle_rotations(a, b).cpu().numpy()
    want = rn.add_axis_angle_rotations(a, b)
    assert rel(got, want) <= 1e-5
    # composition property: R(result) == R(rv1) R(rv2)
    R = rn.rodrigues_batch(got.astype(np.float64))
    assert np.abs(R - rn.rodrigues_batch(a) @ rn.rodrigues_batch(b)).max() < 1e-5


@pytest.mark.parametrize('B', [1, 100, 4096])
def test_rotate_global_pose(B):
    from smplb200 import api
    rng = np.random.default_rng(B)
    th = rng.normal(0, 0.5, size=(B, 72))
    zr = rng.uniform(-np.pi, np.pi, size=B)
    got = api.rotate_global_pose(th, zr).cpu().numpy()
    want = rn.rotate_global_pose(th, zr)
    assert np.array_equal(got[:, 3:], th[:, 3:].astype(np.float32))
    # axis-angle vectors near |c| = pi are ill-conditioned as vectors: compare the rotations they encode
    Rg, Rw = rn.rodrigues_batch(got[:, :3].astype(np.float64)), rn.rodrigues_batch(want[:, :3])
    assert np.abs(Rg - Rw).max() <= 1e-5


def test_surreal_fixture_rotate_global_pose():
    import os
    from conftest import ROOT
    from smplb200 import api
    g = np.load(os.path.join(ROOT, 'tests', 'golden', 'smpl_surreal16.npz'))
    th = g['theta'].astype(np.float64)
    zr = np.linspace(-3, 3, th.shape[0])
    got = api.rotate_global_pose(th, zr).cpu().numpy()
    want = rn.rotate_global_pose(th, zr)
    assert np.abs(rn.rodrigues_batch(got[:, :3].astype(np.float64)) - rn.rodrigues_batch(want[:, :3])).max() <= 1e-5


@pytest.mark.parametrize('B,seed', [(1, 0), (37, 1), (1000, 2)])
def test_cam_losses_value_and_gradient(B, seed):
    from smplb200 import api
    args = cam_inputs(B, seed)
    loss, g = api.cam_losses(args[0], args[1], args[2], args[3], args[4], f_raw=args[5])
    loss = loss.cpu().numpy()
    t = [torch.from_numpy(a).requires_grad_(i in (2, 3, 4, 5)) for i, a in enumerate(args)]
    la = rt.cam_angle_loss(*t[:5])
    ln, lr = rt.cam_reg_losses(t[2], t[3], t[5])
    (la + ln + lr).backward()
    want = np.array([float(la), float(ln), float(lr), float(la + ln + lr)])
    assert np.abs(loss - want).max() <= 1e-5 * np.abs(want).max(), (loss, want)
    for key, i in (('cam_pos', 2), ('cam_rot', 3), ('cam_f', 4), ('f_raw', 5)):
        gw = t[i].grad.numpy()
        assert rel(g[key].cpu().numpy(), gw) <= 1e-4, key


def test_cam_losses_without_gradients_and_f_raw():
    from smplb200 import api
    args = cam_inputs(64, 3)
    loss, g = api.cam_losses(args[0], args[1], args[2], args[3], args[4], want_grad=False)
    assert g is None
    n, r = rn.cam_reg_losses(args[2], args[3], np.zeros(64))
    a = rn.cam_angle_loss(*args[:5])
    want = np.array([a, n, r, a + n + r])
    assert np.abs(loss.cpu().numpy() - want).max() <= 1e-5 * np.abs(want).max()
