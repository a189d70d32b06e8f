"""GPU parity of the hand-written backward (smplb_backward / smplb_project_backward, through the C ABI)
against torch float64 autograd of the oracle twin (what TF autodiff computes, pose_model_3d.py:393-400).
Tolerance (BASELINE.json north_star): max|g - g_ref| / max|g_ref| <= 1e-4."""
import numpy as np
import pytest
import torch

from conftest import maxrel
from oracle import ref_torch as rt
from smplb200 import synthetic as syn, _lib
import smplb200

pytestmark = pytest.mark.gpu
GTOL = 1e-4


def _t64(x, grad=False):
    return torch.tensor(np.asarray(x, dtype=np.float64), requires_grad=grad)


@pytest.fixture(scope='module')
def tmodel64(synth_dd):
    return rt.load_model(synth_dd, dtype=torch.float64)


@pytest.mark.parametrize('B', [1, 5, 64, 131])
def test_joint_path_gradients(gpu_model, tmodel64, B):
    """Linear functionals of J_transformed and of the 2D pixels (camera constant): exercises the projection,
    chain, joint-regression and Rodrigues adjoints."""
    inp = syn.make_inputs(B, seed=300 + B, dtype=np.float32)
    rng = np.random.default_rng(B)
    wj = rng.normal(size=(B, 24, 3)).astype(np.float32)
    wp = rng.normal(size=(B, 24, 2)).astype(np.float32) * 1e-3
    beta, theta = _t64(inp['beta'], True), _t64(inp['theta'], True)
    o = rt.smpl_forward(tmodel64, beta, theta, tile_W=False)
    _, px = rt.project_joints_px(o['J_transformed'], _t64(inp['cam_pos']), _t64(inp['cam_rot']), _t64(inp['focal']))
    loss = (o['J_transformed'] * _t64(wj)).sum() + (px * _t64(wp)).sum()
    loss.backward()
    gb, gt = gpu_model.backward(inp['beta'], inp['theta'], g_joints=wj, g_px=wp, cam_pos=inp['cam_pos'],
                                cam_rot=inp['cam_rot'], focal=inp['focal'])
    assert maxrel(gt.cpu().numpy(), theta.grad.numpy()) <= GTOL
    assert maxrel(gb.cpu().numpy(), beta.grad.numpy()) <= GTOL


def test_reprojection_loss_gradient(gpu_model, tmodel64):
    """The reference's re-projection loss (pose_model_3d.py:218-238, uncapped) end to end:
    forward on the GPU, dL/dpx formed by the caller, backward on the GPU."""
    B = 48
    inp = syn.make_inputs(B, seed=17, dtype=np.float32)
    tgt = syn.make_inputs(B, seed=18, dtype=np.float32)
    cam = [_t64(inp[k]) for k in ('cam_pos', 'cam_rot', 'focal')]
    beta, theta = _t64(inp['beta'], True), _t64(inp['theta'], True)
    with torch.no_grad():
        gt_j = rt.smpl_forward(tmodel64, _t64(inp['beta']), _t64(tgt['theta']), tile_W=False)['J_transformed']
    o = rt.smpl_forward(tmodel64, beta, theta, tile_W=False)
    loss, out_2d, gt_2d = rt.reproj_loss(o['J_transformed'], gt_j, *cam, cap=False)
    loss.backward()
    out = gpu_model.forward(inp['beta'], inp['theta'], inp['cam_pos'], inp['cam_rot'], inp['focal'],
                            flags=_lib.F_NO_SKIN)
    px = out['px']
    assert maxrel(px.cpu().numpy(), out_2d.detach().numpy()) <= 1e-5
    g_px = (2.0 * rt.CFG.ss / (B * 24 * 2)) * (px - torch.tensor(gt_2d.numpy(), dtype=torch.float32, device=px.device))
    gb, gt = gpu_model.backward(inp['beta'], inp['theta'], g_px=g_px, cam_pos=inp['cam_pos'],
                                cam_rot=inp['cam_rot'], focal=inp['focal'])
    assert maxrel(gt.cpu().numpy(), theta.grad.numpy()) <= GTOL
    assert maxrel(gb.cpu().numpy(), beta.grad.numpy()) <= GTOL


def test_zero_theta_gradient_is_finite(gpu_model):
    """theta = 0: eps-in-norm Rodrigues keeps the gradient finite (batch_lbs.py:45-46)."""
    B = 3
    wj = np.ones((B, 24, 3), np.float32)
    gb, gt = gpu_model.backward(np.zeros((B, 10), np.float32), np.zeros((B, 72), np.float32), g_joints=wj)
    assert torch.isfinite(gt).all() and torch.isfinite(gb).all()


@pytest.mark.parametrize('tiled', [True, False])
def test_project_backward(tiled):
    """project() adjoint wrt points and camera (the camera-loss path, pose_model_3d.py:243-245)."""
    rng = np.random.default_rng(4)
    N, P = 24 * 9, 24
    nc = N if tiled else N // P
    pts = rng.normal(0, 0.6, size=(N, 3))
    cp = np.array([0.1, -0.3, -4.0]) + rng.normal(0, 0.05, size=(nc, 3))
    ca = rng.normal(0, 0.1, size=(nc, 3))
    fl = np.full(nc, 0.05)
    w = rng.normal(size=(N, 2))
    tp, tcp, tca, tfl = _t64(pts, True), _t64(cp, True), _t64(ca, True), _t64(fl, True)
    if tiled:
        uv = rt.project(tp, tcp, tca, tfl)
    else:
        uv = rt.project(tp, rt.tile_cam(tcp, P), rt.tile_cam(tca, P), rt.tile_cam(tfl, P))
    (uv * _t64(w)).sum().backward()
    g_pts, g_cp, g_ca, g_fl = smplb200.project_backward(pts.astype(np.float32), cp.astype(np.float32),
                                                        ca.astype(np.float32), fl.astype(np.float32),
                                                        w.astype(np.float32), P=P, cam_tiled=tiled, want_focal=True)
    assert maxrel(g_pts.cpu().numpy(), tp.grad.numpy()) <= GTOL
    assert maxrel(g_cp.cpu().numpy(), tcp.grad.numpy()) <= GTOL
    assert maxrel(g_ca.cpu().numpy(), tca.grad.numpy()) <= GTOL
    assert maxrel(g_fl.cpu().numpy(), tfl.grad.numpy()) <= GTOL          # project.py:35-42 is differentiable in f


def test_dropin_project_autograd():
    """The drop-in project() is differentiable like the reference's (camera loss, pose_model_3d.py:240-254)."""
    from tf_perspective_projection.project import project
    rng = np.random.default_rng(8)
    N = 48
    pts = torch.tensor(rng.normal(0, 0.5, size=(N, 3)), dtype=torch.float32, device='cuda', requires_grad=True)
    cp = torch.tensor(np.tile([[0.0, -0.35, -4.0]], (N, 1)), dtype=torch.float32, device='cuda', requires_grad=True)
    ca = torch.tensor(rng.normal(0, 0.05, size=(N, 3)), dtype=torch.float32, device='cuda', requires_grad=True)
    fl = torch.full((N,), 0.05, device='cuda', requires_grad=True)      # a learnable focal length
    uv = project(pts, cp, ca, fl)
    uv.square().sum().backward()
    tp, tcp, tca = _t64(pts.detach().cpu().numpy(), True), _t64(cp.detach().cpu().numpy(), True), _t64(ca.detach().cpu().numpy(), True)
    tfl = _t64(fl.detach().cpu().numpy(), True)
    rt.project(tp, tcp, tca, tfl).square().sum().backward()
    assert maxrel(pts.grad.cpu().numpy(), tp.grad.numpy()) <= GTOL
    assert maxrel(cp.grad.cpu().numpy(), tcp.grad.numpy()) <= GTOL
    assert maxrel(ca.grad.cpu().numpy(), tca.grad.numpy()) <= GTOL
    assert maxrel(fl.grad.cpu().numpy(), tfl.grad.numpy()) <= GTOL


@pytest.mark.parametrize('B,with_kp', [(3, False), (40, True), (130, False)])
def test_vertex_path_gradients(gpu_model, tmodel64, B, with_kp):
    """Linear functionals of the vertices (+ keypoints, + posed joints): skinning adjoint (tensor-core T^T g and
    the per-joint reduction), transposed blend GEMM, then the joint path (mesh loss, pose_model_3d.py:183-192)."""
    inp = syn.make_inputs(B, seed=900 + B, dtype=np.float32)
    rng = np.random.default_rng(B)
    wv = rng.normal(size=(B, 6890, 3)).astype(np.float32) / 6890
    wj = rng.normal(size=(B, 24, 3)).astype(np.float32) * 0.1
    wk = rng.normal(size=(B, 19, 3)).astype(np.float32) * 0.1 if with_kp else None
    beta, theta = _t64(inp['beta'], True), _t64(inp['theta'], True)
    o = rt.smpl_forward(tmodel64, beta, theta, tile_W=False)
    loss = (o['verts'] * _t64(wv)).sum() + (o['J_transformed'] * _t64(wj)).sum()
    if with_kp:
        loss = loss + (o['joints'] * _t64(wk)).sum()
    loss.backward()
    gpu_model.forward(inp['beta'], inp['theta'], flags=_lib.F_KEEP_VPOSED)
    gb, gt = gpu_model.backward(inp['beta'], inp['theta'], g_verts=wv, g_joints=wj, g_kp=wk)
    assert maxrel(gt.cpu().numpy(), theta.grad.numpy()) <= GTOL
    assert maxrel(gb.cpu().numpy(), beta.grad.numpy()) <= GTOL


def test_dropin_smpl_autograd(synth_dd, tmodel64, tmp_path):
    """The drop-in SMPL class is differentiable (torch.autograd.Function over the hand-written backward):
    mesh + joint MSE as in pose_model_3d.py:183-192."""
    from tf_smpl.batch_smpl import SMPL
    path = str(tmp_path / 'm.pkl')
    syn.write_pkl(path, synth_dd)
    smpl = SMPL(path)
    B = 6
    inp = syn.make_inputs(B, seed=5, dtype=np.float32)
    tgt = syn.make_inputs(B, seed=6, dtype=np.float32)
    with torch.no_grad():
        gt_v, _, _ = smpl(inp['beta'], tgt['theta'], get_skin=True)
        gt_j = smpl.J_transformed.clone()
    beta = torch.tensor(inp['beta'], device='cuda', requires_grad=True)
    theta = torch.tensor(inp['theta'], device='cuda', requires_grad=True)
    verts, _, _ = smpl(beta, theta, get_skin=True)
    loss = 1.0 * torch.mean((verts - gt_v) ** 2) + 2.0 * torch.mean((smpl.J_transformed - gt_j) ** 2)
    loss.backward()
    b64, t64 = _t64(inp['beta'], True), _t64(inp['theta'], True)
    o = rt.smpl_forward(tmodel64, b64, t64, tile_W=False)
    with torch.no_grad():
        og = rt.smpl_forward(tmodel64, _t64(inp['beta']), _t64(tgt['theta']), tile_W=False)
    rt.loss3d(og['verts'], o['verts'], og['J_transformed'], o['J_transformed']).backward()
    assert maxrel(theta.grad.cpu().numpy(), t64.grad.numpy()) <= GTOL
    assert maxrel(beta.grad.cpu().numpy(), b64.grad.numpy()) <= GTOL


def test_dropin_two_calls_then_one_backward(synth_dd, tmodel64, tmp_path):
    """The reference's training step builds TWO SMPL calls before it differentiates (pose_model_3d.py:177-192). Both
    calls keep their own workspace, so one backward through both needs no hidden extra forward: exactly two
    smplb_forward launches, gradients of both calls vs the float64 autograd twin."""
    from tf_smpl.batch_smpl import SMPL
    path = str(tmp_path / 'm.pkl')
    syn.write_pkl(path, synth_dd)
    smpl = SMPL(path)
    B = 5
    a, b = syn.make_inputs(B, seed=11, dtype=np.float32), syn.make_inputs(B, seed=12, dtype=np.float32)
    th1 = torch.tensor(a['theta'], device='cuda', requires_grad=True)
    th2 = torch.tensor(b['theta'], device='cuda', requires_grad=True)
    be = torch.tensor(a['beta'], device='cuda', requires_grad=True)
    n0 = smpl._model.n_forward_calls
    v1, _, _ = smpl(be, th1, get_skin=True)
    j1 = smpl.J_transformed
    v2, _, _ = smpl(be, th2, get_skin=True)          # second call: must not clobber what the first one kept
    j2 = smpl.J_transformed
    loss = 1.0 * torch.mean((v1 - v2) ** 2) + 2.0 * torch.mean((j1 - j2) ** 2)
    loss.backward()
    assert smpl._model.n_forward_calls - n0 == 2      # no re-run of the forward inside backward
    b64, t1, t2 = _t64(a['beta'], True), _t64(a['theta'], True), _t64(b['theta'], True)
    o1 = rt.smpl_forward(tmodel64, b64, t1, tile_W=False)
    o2 = rt.smpl_forward(tmodel64, b64, t2, tile_W=False)
    rt.loss3d(o2['verts'], o1['verts'], o2['J_transformed'], o1['J_transformed']).backward()
    assert maxrel(th1.grad.cpu().numpy(), t1.grad.numpy()) <= GTOL
    assert maxrel(th2.grad.cpu().numpy(), t2.grad.numpy()) <= GTOL
    assert maxrel(be.grad.cpu().numpy(), b64.grad.numpy()) <= GTOL


def test_backward_rejects_a_stale_workspace(gpu_model):
    """Model.backward without an explicit workspace only accepts the cached one when the last forward filled it for
    the same batch with F_KEEP_VPOSED (it used to reallocate silently and read garbage)."""
    from smplb200 import _lib
    inp = syn.make_inputs(4, seed=1, dtype=np.float32)
    gv = np.ones((4, 6890, 3), np.float32)
    gpu_model.forward(inp['beta'], inp['theta'])                       # no F_KEEP_VPOSED
    with pytest.raises(RuntimeError):
        gpu_model.backward(inp['beta'], inp['theta'], g_verts=gv)
    gpu_model.forward(inp['beta'], inp['theta'], flags=_lib.F_KEEP_VPOSED)
    gpu_model.forward(inp['beta'][:2], inp['theta'][:2], flags=_lib.F_KEEP_VPOSED)   # another batch in between
    with pytest.raises(RuntimeError):
        gpu_model.backward(inp['beta'], inp['theta'], g_verts=gv)
    gpu_model.forward(inp['beta'], inp['theta'], flags=_lib.F_KEEP_VPOSED)
    gb, gt = gpu_model.backward(inp['beta'], inp['theta'], g_verts=gv)
    assert torch.isfinite(gt).all()
