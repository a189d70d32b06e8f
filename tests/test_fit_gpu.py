"""Fitting step (smplb_fit_step) against the oracle trajectory: fp64 autograd gradients + the TF1 Adam formula
(pose_model_3d.py:393).  First steps must match to 1e-4."""
import numpy as np
import pytest
import torch

from conftest import maxrel
from oracle import ref_numpy as rn, ref_torch as rt
from smplb200 import synthetic as syn
from smplb200.fit import Fitter

pytestmark = pytest.mark.gpu


def _t64(x, grad=False):
    return torch.tensor(np.asarray(x, dtype=np.float64), requires_grad=grad)


def _oracle_traj(m64, inp, theta0, gt_px, steps, fit_beta, cap, lr):
    beta, theta = inp['beta'].astype(np.float64), theta0.astype(np.float64)
    cam = [_t64(inp[k]) for k in ('cam_pos', 'cam_rot', 'focal')]
    st = {'mt': 0 * theta, 'vt': 0 * theta, 'mb': 0 * beta, 'vb': 0 * beta}
    losses = []
    for t in range(1, steps + 1):
        b, th = _t64(beta, True), _t64(theta, True)
        o = rt.smpl_forward(m64, b, th, tile_W=False)
        _, px = rt.project_joints_px(o['J_transformed'], *cam)
        loss = rt.mse(_t64(gt_px), px, rt.CFG.ss)
        losses.append(float(loss.detach()))
        if cap:
            loss = torch.minimum(torch.tensor(cap, dtype=torch.float64), loss)
        loss.backward()
        theta, st['mt'], st['vt'] = rn.adam_tf1(theta, th.grad.numpy(), st['mt'], st['vt'], t, lr=lr)
        if fit_beta:
            beta, st['mb'], st['vb'] = rn.adam_tf1(beta, b.grad.numpy(), st['mb'], st['vb'], t, lr=lr)
    return beta, theta, losses


@pytest.mark.parametrize('fit_beta', [False, True])
def test_first_steps_match_oracle(gpu_model, synth_dd, fit_beta):
    B, steps, lr = 24, 3, 1e-2
    m64 = rt.load_model(synth_dd, dtype=torch.float64)
    inp = syn.make_inputs(B, seed=21, dtype=np.float32)
    rng = np.random.default_rng(3)
    theta0 = (inp['theta'] + rng.normal(0, 0.1, size=inp['theta'].shape)).astype(np.float32)
    o = rn.smpl_forward(rn.load_model(synth_dd, dtype=np.float64), inp['beta'].astype(np.float64),
                        inp['theta'].astype(np.float64), tile_W=False)
    _, gt_px = rn.project_joints_px(o['J_transformed'], *(inp[k].astype(np.float64) for k in ('cam_pos', 'cam_rot', 'focal')))
    gt_px = gt_px.astype(np.float32)
    f = Fitter(gpu_model, inp['beta'], theta0, inp['cam_pos'], inp['cam_rot'], gt_px, focal=inp['focal'],
               fit_beta=fit_beta, lr=lr)
    losses = [float(f.step()[0]) for _ in range(steps)]
    beta_ref, theta_ref, loss_ref = _oracle_traj(m64, inp, theta0, gt_px.astype(np.float64), steps, fit_beta, 0.0, lr)
    assert maxrel(np.array(losses), np.array(loss_ref)) <= 1e-4
    assert maxrel(f.theta.cpu().numpy() - theta0, theta_ref - theta0) <= 1e-3   # 3 Adam steps of size lr
    assert maxrel(f.theta.cpu().numpy(), theta_ref) <= 1e-4
    assert maxrel(f.beta.cpu().numpy(), beta_ref) <= 1e-4
    assert losses[-1] < losses[0]


def test_reference_cap_gates_the_gradient(gpu_model):
    """pose_model_3d.py:237: min(1/24, loss) -> zero gradient while the scaled MSE exceeds the cap."""
    B = 8
    inp = syn.make_inputs(B, seed=2, dtype=np.float32)
    gt_px = np.zeros((B, 24, 2), np.float32)          # far away: loss >> 1/24
    f = Fitter(gpu_model, inp['beta'], inp['theta'], inp['cam_pos'], inp['cam_rot'], gt_px, reproj_cap=1.0 / 24)
    l = f.step().cpu().numpy()
    assert l[0] > 1.0 / 24 and abs(l[1] - 1.0 / 24) < 1e-7 and l[2] == 0.0
    assert np.array_equal(f.theta.cpu().numpy(), inp['theta'])           # no update under the cap


def test_fit_step_in_cuda_graph(gpu_model, synth_dd):
    """The step is allocation-free and stream-ordered: replaying it from a CUDA graph gives the same result."""
    B = 64
    inp = syn.make_inputs(B, seed=9, dtype=np.float32)
    tgt = syn.make_inputs(B, seed=10, dtype=np.float32)
    out = gpu_model.forward(inp['beta'], tgt['theta'], inp['cam_pos'], inp['cam_rot'], inp['focal'], flags=4)
    gt_px = out['px'].clone()
    fa = Fitter(gpu_model, inp['beta'], inp['theta'], inp['cam_pos'], inp['cam_rot'], gt_px, lr=1e-2)
    fb = Fitter(gpu_model, inp['beta'], inp['theta'], inp['cam_pos'], inp['cam_rot'], gt_px, lr=1e-2)
    fa.step()
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            fb.step()                                     # captured with step = 1 (bias correction of step 1)
    fb.theta.copy_(torch.as_tensor(inp['theta']).cuda()); fb.m_theta.zero_(); fb.v_theta.zero_()
    g.replay()
    torch.cuda.synchronize()
    assert torch.equal(fa.theta, fb.theta)


def test_mse_grad_value_and_gradient():
    """smplb_mse_grad = tf.losses.mean_squared_error with a scalar weight (pose_model_3d.py:183-192)."""
    import smplb200
    rng = np.random.default_rng(0)
    for n, w in ((1, 1.0), (1000, 2.0), (3 * 6890 * 5, 1.0)):
        a, b = rng.normal(size=n).astype(np.float32), rng.normal(size=n).astype(np.float32)
        loss, g = smplb200.mse_grad(a, b, weight=w)
        want = w * np.mean((a.astype(np.float64) - b) ** 2)
        assert abs(float(loss[0]) - want) <= 1e-5 * want
        assert maxrel(g.cpu().numpy(), 2 * w / n * (a.astype(np.float64) - b)) <= 1e-6
        loss2, _ = smplb200.mse_grad(a, b, weight=w, loss_out=loss, accumulate=True, want_grad=False)
        assert abs(float(loss2[0]) - 2 * want) <= 1e-5 * want
        l3, _ = smplb200.mse_grad(a, b, weight=w, want_grad=False)
        assert float(l3[0]) == float(smplb200.mse_grad(a, b, weight=w, want_grad=False)[0][0])   # run-to-run identical


def test_mesh_fit_first_steps_match_oracle(gpu_model, synth_dd):
    """3D-loss fitting (mesh + joint MSE, pose_model_3d.py:183-192) through the vertex backward vs the float64
    autograd + TF1-Adam trajectory."""
    from smplb200.fit import MeshFitter
    B, steps, lr = 6, 3, 1e-2
    m64 = rt.load_model(synth_dd, dtype=torch.float64)
    inp = syn.make_inputs(B, seed=31, dtype=np.float32)
    rng = np.random.default_rng(4)
    theta0 = (inp['theta'] + rng.normal(0, 0.1, size=inp['theta'].shape)).astype(np.float32)
    with torch.no_grad():
        tgt = rt.smpl_forward(m64, _t64(inp['beta']), _t64(inp['theta']), tile_W=False)
    gv, gj = tgt['verts'].numpy(), tgt['J_transformed'].numpy()
    f = MeshFitter(gpu_model, inp['beta'], theta0, gv.astype(np.float32), gj.astype(np.float32), lr=lr)
    losses = [float(f.step()[0]) for _ in range(steps)]
    theta = theta0.astype(np.float64)
    mt, vt, want = 0 * theta, 0 * theta, []
    for t in range(1, steps + 1):
        th = _t64(theta, True)
        o = rt.smpl_forward(m64, _t64(inp['beta']), th, tile_W=False)
        loss = rt.loss3d(_t64(gv), o['verts'], _t64(gj), o['J_transformed'])
        want.append(float(loss.detach()))
        loss.backward()
        theta, mt, vt = rn.adam_tf1(theta, th.grad.numpy(), mt, vt, t, lr=lr)
    assert maxrel(np.array(losses), np.array(want)) <= 1e-4
    assert maxrel(f.theta.cpu().numpy(), theta) <= 1e-4
    assert losses[-1] < losses[0]
