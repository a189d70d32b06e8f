"""Kinematic trees other than SMPL's through the per-sample kernels (pose_prep_kernel, joint_bwd_kernel): the forward
chain runs by pointer jumping (ceil(log2(depth + 1)) rounds) and the chain adjoint as range sums over the joints in
depth-first order, both driven by tables built from `parents` (csrc/model.cu build_tree) -- so a 24-joint chain (depth 23,
5 rounds), a ternary tree and random trees must match the reference recursion (batch_lbs.py:122-136) exactly like SMPL's
own tree does.  Oracle: float64 restatement + autograd.  Tolerances as in test_forward_gpu / test_backward_gpu."""
import numpy as np
import pytest
import torch

from conftest import maxrel
from oracle import ref_torch as rt
from smplb200 import synthetic as syn
import smplb200

pytestmark = pytest.mark.gpu


def _trees():
    out = {'chain': np.arange(-1, 23), 'ternary': np.array([-1] + [(i - 1) // 3 for i in range(1, 24)])}
    rng = np.random.default_rng(11)
    for k in range(2):
        par, nch = [-1], [0] * 24
        for i in range(1, 24):
            while True:
                p = int(rng.integers(0, i))
                if nch[p] < 3:       # the C ABI supports at most 3 children per joint
                    break
            par.append(p)
            nch[p] += 1
        out['random%d' % k] = np.array(par)
    return out


TREES = _trees()


def _t64(x, grad=False):
    return torch.tensor(np.asarray(x, dtype=np.float64), requires_grad=grad)


@pytest.mark.parametrize('name', sorted(TREES))
def test_other_trees_forward_and_gradients(name):
    dd = dict(syn.make_synthetic_smpl(0))
    kt = np.array(dd['kintree_table']).copy()
    kt[0, 1:] = TREES[name][1:].astype(kt.dtype)
    dd['kintree_table'] = kt
    model = smplb200.Model(dd, device=0)
    m64 = rt.load_model(dd, dtype=torch.float64)
    B = 37
    inp = syn.make_inputs(B, seed=5, dtype=np.float32)
    inp['theta'] = (inp['theta'] * 0.3).astype(np.float32)     # a 23-deep chain multiplies 24 rotations
    rng = np.random.default_rng(1)
    wj = rng.normal(size=(B, 24, 3)).astype(np.float32)
    wp = rng.normal(size=(B, 24, 2)).astype(np.float32) * 1e-3
    wv = rng.normal(size=(B, 6890, 3)).astype(np.float32) * 1e-2
    beta, theta = _t64(inp['beta'], True), _t64(inp['theta'], True)
    o = rt.smpl_forward(m64, beta, theta, tile_W=False)
    cam = [_t64(inp[k]) for k in ('cam_pos', 'cam_rot', 'focal')]
    _, px = rt.project_joints_px(o['J_transformed'], *cam)
    got = model.forward(inp['beta'], inp['theta'], inp['cam_pos'], inp['cam_rot'], inp['focal'])
    scale = float(o['verts'].detach().abs().max())
    assert float(np.abs(got['verts'].cpu().numpy() - o['verts'].detach().numpy()).max()) <= 1e-5 * max(scale, 1.0)
    assert maxrel(got['joints'].cpu().numpy(), o['J_transformed'].detach().numpy()) <= 1e-5
    assert maxrel(got['px'].cpu().numpy(), px.detach().numpy()) <= 1e-5
    # joints-only adjoint (the fitting path) and the vertex path
    ((o['J_transformed'] * _t64(wj)).sum() + (px * _t64(wp)).sum()).backward()
    gb, gt = model.backward(inp['beta'], inp['theta'], g_joints=wj, g_px=wp, cam_pos=inp['cam_pos'], cam_rot=inp['cam_rot'],
                            focal=inp['focal'])
    assert maxrel(gt.cpu().numpy(), theta.grad.numpy()) <= 1e-4
    assert maxrel(gb.cpu().numpy(), beta.grad.numpy()) <= 1e-4
    # leaves of the tree get an exact zero from the joints-only loss, as TF's autodiff gives
    leaves = [j for j in range(24) if j not in set(TREES[name][1:].tolist())]
    g3 = gt.cpu().numpy().reshape(B, 24, 3)
    assert np.all(g3[:, leaves] == 0.0)
    beta2, theta2 = _t64(inp['beta'], True), _t64(inp['theta'], True)
    o2 = rt.smpl_forward(m64, beta2, theta2, tile_W=False)
    (o2['verts'] * _t64(wv)).sum().backward()
    model.forward(inp['beta'], inp['theta'], flags=smplb200._lib.F_KEEP_VPOSED)
    gb2, gt2 = model.backward(inp['beta'], inp['theta'], g_verts=wv)
    assert maxrel(gt2.cpu().numpy(), theta2.grad.numpy()) <= 1e-4
    assert maxrel(gb2.cpu().numpy(), beta2.grad.numpy()) <= 1e-4
