"""CPU: the N2/N3 oracle restatements (oracle/ref_numpy.py, ref_torch.py): numpy vs the torch twin, analytic
identities, autograd vs finite differences for the camera losses (pose_model_3d.py:256-292)."""
import numpy as np
import torch

from oracle import ref_numpy as rn, ref_torch as rt


def cam_inputs(B, seed, dtype=np.float64):
    rng = np.random.default_rng(seed)
    j3 = rng.normal(0, 0.4, size=(B, 24, 3))
    pos = np.stack([rng.normal(0, 0.1, B), rng.normal(-0.35, 0.1, B), rng.uniform(-6, -3, B)], 1)
    rot = rng.normal(0, 0.25, size=(B, 3))           # some norms above pi/8, some below
    f = np.full(B, 0.05)
    _, px = rn.project_joints_px(j3, pos, rot, f)
    p2d = px + rng.normal(0, 6.0, size=px.shape)     # noisy 2D ground truth
    f_raw = rng.normal(0.02, 0.05, B)
    pos[0, 2] = 0.7                                   # violates -z - 0.3 >= 0
    pos[1 % B, 2] = -11.0                             # violates z + 10 >= 0
    return [x.astype(dtype) for x in (j3, p2d, pos, rot, f, f_raw)]


def test_angle_loss_zero_for_exact_projection():
    j3, _, pos, rot, f, _ = cam_inputs(6, 0)
    pos[:, 2] = -4.0                                  # every joint in front of the camera
    _, px = rn.project_joints_px(j3, pos, rot, f)
    # rays through the exact projections coincide with the rays to the 3D joints: angle == clip floor
    loss = rn.cam_angle_loss(j3, px, pos, rot, f)
    assert loss <= 10.0 * np.arccos(1 - 1e-6) ** 2 * 1.01


def test_numpy_matches_torch_twin():
    args = cam_inputs(9, 1)
    t = [torch.from_numpy(a) for a in args]
    assert abs(rn.cam_angle_loss(*args[:5]) - float(rt.cam_angle_loss(*t[:5]))) < 1e-12
    n, r = rn.cam_reg_losses(args[2], args[3], args[5])
    tn, tr = rt.cam_reg_losses(t[2], t[3], t[5])
    assert abs(n - float(tn)) < 1e-12 and abs(r - float(tr)) < 1e-12
    assert n > 0 and r > 0


def test_autograd_vs_finite_differences():
    args = cam_inputs(5, 2)
    t = [torch.from_numpy(a).requires_grad_(i in (2, 3)) for i, a in enumerate(args)]
    L = rt.cam_angle_loss(*t[:5]) + sum(rt.cam_reg_losses(t[2], t[3], t[5]))
    L.backward()

    def f(pos, rot):
        return rn.cam_angle_loss(args[0], args[1], pos, rot, args[4]) + sum(rn.cam_reg_losses(pos, rot, args[5]))

    rng = np.random.default_rng(0)
    for which in (2, 3):
        g = t[which].grad.numpy()
        for _ in range(6):
            i, c = rng.integers(0, 5), rng.integers(0, 3)
            d = np.zeros_like(args[which])
            d[i, c] = 1e-6
            pp = (args[2] + d, args[3]) if which == 2 else (args[2], args[3] + d)
            pm = (args[2] - d, args[3]) if which == 2 else (args[2], args[3] - d)
            fd = (f(*pp) - f(*pm)) / 2e-6
            assert abs(fd - g[i, c]) <= 1e-5 * max(1.0, abs(fd))


def test_rotate_global_pose_zero_rotation_branches():
    th = np.zeros((3, 72))
    th[1, :3] = [0.3, -0.2, 0.5]
    out = rn.rotate_global_pose(th, np.array([0.0, 0.4, -1.0]))
    assert np.isfinite(out).all()
    R = rn.rodrigues_batch(out[:, :3])
    for b in range(3):
        assert np.allclose(R[b] @ R[b].T, np.eye(3), atol=1e-12)
