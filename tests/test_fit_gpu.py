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
    """The step is allocation-free and stream-ordered, and Adam's step counter lives on the device: N replays of ONE
    captured step equal N eager steps (the bias correction lr_t advances on every replay, pose_model_3d.py:393)."""
    B, N = 64, 6
    inp = syn.make_inputs(B, seed=9, dtype=np.float32)
    tgt = syn.make_inputs(B, seed=10, dtype=np.float32)
    out = gpu_model.forward(inp['beta'], tgt['theta'], inp['cam_pos'], inp['cam_rot'], inp['focal'], flags=4)
    gt_px = out['px'].clone()
    fa = Fitter(gpu_model, inp['beta'], inp['theta'], inp['cam_pos'], inp['cam_rot'], gt_px, lr=1e-2)
    fb = Fitter(gpu_model, inp['beta'], inp['theta'], inp['cam_pos'], inp['cam_rot'], gt_px, lr=1e-2)
    la = [fa.step().clone() for _ in range(N)]
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            fb.step()                                     # captured once; nothing runs during capture
    torch.cuda.current_stream().wait_stream(s)
    lb = []
    for _ in range(N):
        g.replay()
        lb.append(fb.loss.clone())
    torch.cuda.synchronize()
    assert int(fb.step_dev) == N and int(fa.step_dev) == N
    assert torch.equal(fa.theta, fb.theta)
    assert torch.equal(torch.stack(la), torch.stack(lb))
    # lr_t of step N against the formula (TF1: lr sqrt(1 - b2^t) / (1 - b1^t))
    want = 1e-2 * np.sqrt(1 - 0.999 ** N) / (1 - 0.9 ** N)
    assert abs(float(lb[-1][3]) - want) <= 1e-5 * want
    # a frozen counter (the round-1 behaviour) would differ: lr_t(1) != lr_t(N)
    assert abs(float(lb[0][3]) - float(lb[-1][3])) > 1e-3 * want


def test_cap_with_sharded_mean_is_rejected(gpu_model):
    """The cap of pose_model_3d.py:237 gates on the GLOBAL loss; a shard cannot decide it."""
    from smplb200._lib import SmplbError, E_UNSUPPORTED
    B = 8
    inp = syn.make_inputs(B, seed=2, dtype=np.float32)
    gt_px = np.zeros((B, 24, 2), np.float32)
    f = Fitter(gpu_model, inp['beta'], inp['theta'], inp['cam_pos'], inp['cam_rot'], gt_px, reproj_cap=1.0 / 24,
               n_total=2 * B * 48)
    with pytest.raises(SmplbError) as e:
        f.step()
    assert e.value.code == E_UNSUPPORTED


def test_mesh_fit_in_cuda_graph(gpu_model, synth_dd):
    """MeshFitter (forward -> mse_grad x2 -> vertex backward -> Adam with the device step counter) replayed from one
    CUDA graph equals the eager steps."""
    from smplb200.fit import MeshFitter
    B, N = 40, 4
    inp = syn.make_inputs(B, seed=31, dtype=np.float32)
    tgt = syn.make_inputs(B, seed=32, dtype=np.float32)
    o = gpu_model.forward(inp['beta'], tgt['theta'], want_kp=False, want_Rs=False)
    gv, gj = o['verts'].clone(), o['joints'].clone()
    fa = MeshFitter(gpu_model, inp['beta'], inp['theta'], gv, gj, lr=1e-2)
    fb = MeshFitter(gpu_model, inp['beta'], inp['theta'], gv, gj, lr=1e-2)
    fb.step()                                             # allocates fb's buffers outside the capture
    fb.theta.copy_(fa.theta); fb.m_theta.zero_(); fb.v_theta.zero_(); fb.step_dev.zero_()
    for _ in range(N):
        fa.step()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            fb.step()
    torch.cuda.current_stream().wait_stream(s)
    for _ in range(N):
        g.replay()
    torch.cuda.synchronize()
    assert int(fb.step_dev) == N
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


def test_baseline_config3_B4096_first_steps(gpu_model, synth_dd):
    """BASELINE.json configs[2] at exactly B = 4096: the first 3 Adam steps of the re-projection fit against the float64
    oracle trajectory (autograd gradients + TF1 Adam), oracle evaluated in chunks (samples are independent; only the
    mean couples them)."""
    B, steps, lr, chunk = 4096, 3, 1e-2, 512
    m64 = rt.load_model(synth_dd, dtype=torch.float64)
    inp = syn.make_inputs(B, seed=4096, dtype=np.float32)
    tgt = syn.make_inputs(B, seed=4097, dtype=np.float32)
    gt_px = gpu_model.forward(inp['beta'], tgt['theta'], inp['cam_pos'], inp['cam_rot'], inp['focal'], flags=4)['px'].cpu().numpy()
    f = Fitter(gpu_model, inp['beta'], inp['theta'], inp['cam_pos'], inp['cam_rot'], gt_px, focal=inp['focal'], lr=lr)
    losses = [float(f.step()[0]) for _ in range(steps)]
    theta = inp['theta'].astype(np.float64)
    mt, vt = 0 * theta, 0 * theta
    want = []
    n_total = B * 48
    for t in range(1, steps + 1):
        g = np.zeros_like(theta)
        tot = 0.0
        for b0 in range(0, B, chunk):
            sl = slice(b0, b0 + chunk)
            th = _t64(theta[sl], True)
            o = rt.smpl_forward(m64, _t64(inp['beta'][sl]), th, tile_W=False, want_verts=False)
            _, px = rt.project_joints_px(o['J_transformed'], *[_t64(inp[k][sl]) for k in ('cam_pos', 'cam_rot', 'focal')])
            part = ((px - _t64(gt_px[sl])) ** 2).sum() * (rt.CFG.ss / n_total)
            part.backward()
            g[sl] = th.grad.numpy()
            tot += float(part.detach())
        want.append(tot)
        theta, mt, vt = rn.adam_tf1(theta, g, mt, vt, t, lr=lr)
    assert maxrel(np.array(losses), np.array(want)) <= 1e-4
    assert maxrel(f.theta.cpu().numpy(), theta) <= 1e-4
    assert maxrel(f.theta.cpu().numpy() - inp['theta'], theta - inp['theta']) <= 1e-3
