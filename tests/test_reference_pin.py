"""Pins the oracle restatements (oracle/ref_numpy.py, ref_torch.py, ref_render.py) on what the REFERENCE'S OWN PYTHON
FILES compute: tests/golden/reference_*.npz were produced by tests/golden/make_reference_golden.py, which imports the
reference's modules from /root/reference and executes them on oracle/tfshim (torch-CPU stand-in for `tensorflow`).

CPU only.  The fixture comparison runs everywhere; when the reference tree is present (the build container) the
reference is also re-run live and must reproduce the committed fixtures."""
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, '3d-pose-estimation_b200'))

from oracle import ref_numpy as rn  # noqa: E402
from oracle import ref_torch as rt  # noqa: E402
from oracle import run_reference as rr  # noqa: E402
from smplb200 import synthetic as syn  # noqa: E402

G = os.path.join(ROOT, 'tests', 'golden')
TIGHT = 1e-12      # float64 restatement vs float64 reference: summation order only


def rel(a, b):
    return float(np.abs(np.asarray(a) - np.asarray(b)).max() / max(np.abs(np.asarray(b)).max(), 1e-300))


@pytest.fixture(scope='module')
def model64():
    return rn.load_model(syn.make_synthetic_smpl(0), dtype=np.float64)


@pytest.fixture(scope='module')
def smpl_g():
    return np.load(os.path.join(G, 'reference_smpl.npz'))


def test_smpl_forward_matches_reference_source(model64, smpl_g):
    g = smpl_g
    f64 = {k: g[k].astype(np.float64) for k in ('beta', 'theta', 'cam_pos', 'cam_rot', 'focal')}
    o = rn.smpl_forward(model64, f64['beta'], f64['theta'], tile_W=True)
    vs = g['vsub']
    assert rel(o['verts'][:, vs], g['verts_sub']) < TIGHT
    assert rel(o['verts'].sum(1), g['verts_sum']) < 1e-11
    assert rel(o['joints'], g['joints']) < TIGHT
    assert rel(o['J_transformed'], g['J_transformed']) < TIGHT
    assert rel(o['Rs'], g['Rs']) < TIGHT
    uv, px = rn.project_joints_px(o['J_transformed'], f64['cam_pos'], f64['cam_rot'], f64['focal'])
    assert rel(uv, g['uv']) < TIGHT
    assert rel(px, g['px']) < TIGHT
    # the fused (no tile(W), no 4x4 homogeneous product) formulation the CUDA path uses is the same function
    o2 = rn.smpl_forward(model64, f64['beta'], f64['theta'], tile_W=False)
    assert rel(o2['verts'], o['verts']) < TIGHT


def test_reference_float32_noise_floor(smpl_g):
    """What the reference itself loses by running in float32 (its native width): the 1e-5 parity bound of the CUDA
    path is set against the float64 values, and the reference's own float32 run must sit inside the same bound."""
    g = smpl_g
    assert rel(g['f32_verts_sub'], g['verts_sub']) < 1e-5
    assert rel(g['f32_J_transformed'], g['J_transformed']) < 1e-5
    assert rel(g['f32_joints'], g['joints']) < 1e-5
    assert rel(g['f32_px'], g['px']) < 1e-5


def _cotangents(B, seed):
    r = np.random.default_rng(seed)
    return (r.standard_normal((B, 6890, 3)) / 6890.0, r.standard_normal((B, 24, 3)), r.standard_normal((B, 19, 3)))


def test_torch_twin_gradients_match_tf_gradients(smpl_g):
    """ref_torch (autograd, float64) is the gradient oracle of the CUDA backward; tf.gradients of the reference graph
    (through the stand-in) must give the same numbers."""
    g = smpl_g
    m = rt.load_model(syn.make_synthetic_smpl(0), dtype=torch.float64)
    B = g['beta'].shape[0]
    beta = torch.tensor(g['beta'].astype(np.float64), requires_grad=True)
    theta = torch.tensor(g['theta'].astype(np.float64), requires_grad=True)
    o = rt.smpl_forward(m, beta, theta, tile_W=False)
    gv, gj, gk = (torch.from_numpy(x) for x in _cotangents(B, 4321))
    L = (o['verts'] * gv).sum() + (o['J_transformed'] * gj).sum() + (o['joints'] * gk).sum()
    gb, gt = torch.autograd.grad(L, [beta, theta], retain_graph=True)
    assert abs(float(L) - float(g['L'])) < 1e-10 * abs(float(g['L']))
    assert rel(gb.numpy(), g['g_beta']) < 1e-11
    assert rel(gt.numpy(), g['g_theta']) < 1e-11
    _, px = rt.project_joints_px(o['J_transformed'], torch.tensor(g['cam_pos'].astype(np.float64)),
                                 torch.tensor(g['cam_rot'].astype(np.float64)), torch.tensor(g['focal'].astype(np.float64)))
    Lp = (px * torch.from_numpy(g['g_px'])).sum()
    gb, gt = torch.autograd.grad(Lp, [beta, theta])
    assert rel(gb.numpy(), g['gpx_beta']) < 1e-10
    assert rel(gt.numpy(), g['gpx_theta']) < 1e-10


@pytest.mark.parametrize('case', ['near', 'far'])
def test_loss_terms_match_get_encoder_losses(model64, case):
    """Every term of PoseModel3d.get_encoder_losses (pose_model_3d.py:142-334), restated term by term."""
    g = np.load(os.path.join(G, 'reference_losses.npz'))
    p = lambda k: g[case + '_' + k]
    outputs, gt_pose, betas, gt2d = p('outputs'), p('gt_pose'), p('betas'), p('gt_joints2d')
    out_pose = outputs[:, :72]
    B = outputs.shape[0]
    cam_pos, cam_rot, f_raw = outputs[:, 72:75], outputs[:, 75:78], outputs[:, 78]
    focal = np.full(B, rn.CFG.fl)                                                     # :212 (constant focal length)
    out_gt_global = np.concatenate([gt_pose[:, :3], out_pose[:, 3:]], axis=1)         # :175-176
    o = rn.smpl_forward(model64, betas, out_gt_global)
    q = rn.smpl_forward(model64, betas, gt_pose)
    l3d = rn.loss3d(q['verts'], o['verts'], q['J_transformed'], o['J_transformed'])
    assert abs(l3d - p('term_3d_loss')) < TIGHT * max(1.0, abs(l3d)) + 1e-15
    rl, _, _ = rn.reproj_loss(o['J_transformed'], q['J_transformed'], cam_pos, cam_rot, focal, cap=False)
    assert abs(rl - p('term_reproj_loss')) < 1e-11 * abs(rl)
    flipped = rn.flip_gt2d(gt2d)
    cl, _ = rn.cam_loss(q['J_transformed'], flipped, cam_pos, cam_rot, focal, cap=False)
    assert abs(cl - p('term_cam_loss')) < 1e-11 * abs(cl)
    al = rn.cam_angle_loss(q['J_transformed'], flipped, cam_pos, cam_rot, focal)
    assert abs(al - p('term_cam_angle')) < 1e-9 * abs(al)
    neg, rot = rn.cam_reg_losses(cam_pos, cam_rot, f_raw)
    pose_mat = rn.mse(rn.rodrigues_batch(gt_pose.reshape(-1, 3)), rn.rodrigues_batch(out_pose.reshape(-1, 3)), 1.0)
    assert abs(pose_mat - p('term_pose_loss')) < 1e-11 * abs(pose_mat)                # :146-159
    total = (pose_mat + float(p('term_reg_loss')) + l3d + min(1.0 / 24, rl) + min(0.5, cl) + al + neg + rot)
    assert abs(total - float(p('total'))) < 1e-10 * abs(total)


@pytest.mark.parametrize('case', ['near', 'far'])
def test_loss_gradients_match_tf_gradients(case):
    """Gradients of the loss terms w.r.t. the network output vector: torch twin vs tf.gradients of the reference."""
    g = np.load(os.path.join(G, 'reference_losses.npz'))
    p = lambda k: g[case + '_' + k]
    m = rt.load_model(syn.make_synthetic_smpl(0), dtype=torch.float64)
    outputs = torch.tensor(p('outputs'), requires_grad=True)
    gt_pose, betas = torch.tensor(p('gt_pose')), torch.tensor(p('betas'))
    gt2d = torch.tensor(p('gt_joints2d'))
    B = outputs.shape[0]
    out_pose = outputs[:, :72]
    cam_pos, cam_rot, f_raw = outputs[:, 72:75], outputs[:, 75:78], outputs[:, 78]
    focal = torch.full((B,), rn.CFG.fl, dtype=torch.float64)
    o = rt.smpl_forward(m, betas, torch.cat([gt_pose[:, :3], out_pose[:, 3:]], 1), tile_W=False)
    q = rt.smpl_forward(m, betas, gt_pose, tile_W=False)
    terms = {}
    terms['3d_loss'] = rt.loss3d(q['verts'], o['verts'], q['J_transformed'], o['J_transformed'])
    terms['reproj_loss'] = rt.reproj_loss(o['J_transformed'], q['J_transformed'], cam_pos, cam_rot, focal, cap=False)[0]
    flipped = gt2d * torch.tensor([1.0, -1.0], dtype=torch.float64) + torch.tensor([0.0, 240.0], dtype=torch.float64)
    terms['cam_loss'] = rt.cam_loss(q['J_transformed'], flipped, cam_pos, cam_rot, focal, cap=False)[0]
    terms['cam_angle'] = rt.cam_angle_loss(q['J_transformed'], flipped, cam_pos, cam_rot, focal)
    for k, v in terms.items():
        gr = torch.autograd.grad(v, outputs, retain_graph=True)[0].numpy()
        ref = p('g_' + k)
        assert abs(float(v) - float(p('term_' + k))) < 1e-9 * abs(float(v)), k
        assert rel(gr, ref) < 1e-8, k
        if k == 'reproj_loss':
            assert np.abs(ref[:, 72:]).max() == 0.0      # camera is stop_gradient'ed (:218-221)
            assert np.abs(ref[:, :3]).max() == 0.0       # global rotation comes from the ground truth (:175-176)
    neg, rot = rt.cam_reg_losses(cam_pos, cam_rot, f_raw)
    pose_mat = rt.mse(rt.rodrigues_batch(gt_pose.reshape(-1, 3)), rt.rodrigues_batch(out_pose.reshape(-1, 3)))
    cap = lambda x, c: torch.minimum(torch.tensor(c, dtype=torch.float64), x)
    total = (pose_mat + terms['3d_loss'] + cap(terms['reproj_loss'], 1.0 / 24) + cap(terms['cam_loss'], 0.5) +
             terms['cam_angle'] + neg + rot)        # reg_loss is identically 0 here (no |joint angle| > 2 pi)
    gr = torch.autograd.grad(total, outputs)[0].numpy()
    assert rel(gr, p('g_total')) < 1e-8


def test_pose_utils_match_reference_source():
    g = np.load(os.path.join(G, 'reference_utils.npz'))
    assert rel(rn.add_axis_angle_rotations(g['rv1'], g['rv2']), g['add_axis_angle']) < 1e-11
    assert rel(rn.rotate_global_pose(g['rgp_thetas'], g['rgp_zrot']), g['rotate_global_pose']) < 1e-11
    assert rel(rn.get_2d_points_in_3d(g['p2d'], g['p2d_cam_rot'], g['p2d_cam_f']), g['points_2d_in_3d']) < TIGHT
    assert rel(rn.batch_orth_proj_idrot(g['orth_X'], g['orth_cam']), g['orth_proj']) < TIGHT
    parents = syn.make_synthetic_smpl(0)['kintree_table'][0].astype(np.int32)
    for rb in (0, 1):
        nj, A = rn.batch_global_rigid_transformation(g['grt_Rs'], g['grt_Js'], parents, rotate_base=bool(rb))
        assert rel(nj, g['grt_newJ_%d' % rb]) < TIGHT
        assert rel(A, g['grt_A_%d' % rb]) < TIGHT
    Rs = rn.batch_rodrigues(g['lrot_theta'][:, 3:].reshape(-1, 3)).reshape(-1, 23, 3, 3)
    assert rel((Rs - np.eye(3)).reshape(-1, 207), g['lrotmin']) < TIGHT               # batch_lbs.py:60-81


def test_normals_match_reference_source():
    g = np.load(os.path.join(G, 'reference_utils.npz'))
    faces = g['faces'].astype(np.int64)
    vf = rn.vertex_faces_from_face_verts(faces)
    assert np.array_equal(vf, g['vertex_faces'].astype(np.int64))
    m = rn.load_model(syn.make_synthetic_smpl(0), dtype=np.float64)
    verts = rn.smpl_forward(m, g['mesh_beta'].astype(np.float64), g['mesh_theta'].astype(np.float64))['verts']
    assert rel(verts, g['mesh_verts']) < 1e-6          # stored as float32
    n = rn.normals_from_mesh(verts, faces, vf)
    assert rel(n, g['normals']) < 1e-11


def test_render_graph_matches_reference_source():
    """utils.render_mesh_verts_cam (utils.py:12-58) -> mesh_renderer.py -> rasterize_triangles.py -> the C++ op:
    the oracle restatement of that whole graph (ref_numpy + ref_render) against the reference's own run (float32)."""
    from oracle import raster as orc
    from oracle import ref_render as rr_
    if not (orc.ref_available() or os.path.exists(os.path.join(ROOT, 'oracle', 'libraster_port.so'))):
        pytest.skip('oracle rasteriser not built')
    g = np.load(os.path.join(G, 'reference_utils.npz'))
    faces = g['faces'].astype(np.int32)
    vf = g['vertex_faces'].astype(np.int64)
    verts = g['mesh_verts'][:1]
    B = 1
    cam_pos, cam_rot = g['render_cam_pos'][:1], g['render_cam_rot'][:1]
    R = rn.rodrigues_batch(cam_rot)
    lookat = (R @ np.array([0.0, 0.0, 1.0], dtype=np.float32)) + cam_pos               # utils.py:22-28
    up = R @ np.array([0.0, 1.0, 0.0], dtype=np.float32)                               # :30-34
    normals = rn.normals_from_mesh(verts, faces, vf).astype(np.float32)
    lights = cam_pos[:, None, :]
    img = rr_.mesh_renderer(verts, faces, normals, np.ones_like(verts), cam_pos, lookat, up, lights,
                            np.ones_like(lights), 320, 240, fov_y=g['render_fov'][:1], near_clip=np.float32(rn.CFG.fl),
                            far_clip=np.float32(100.0))
    lit = img[..., :3].mean(axis=3, keepdims=True)                                     # :55
    ref = g['render_lit'][:1]
    assert lit.shape == ref.shape == (B, 240, 320, 1)
    diff = np.abs(lit - ref)
    assert (diff > 1e-4).mean() < 1e-4, ((diff > 1e-4).mean(), diff.max())             # edge pixels may flip in fp32
    sil = rr_.mesh_renderer(verts, faces, np.zeros_like(verts), np.ones_like(verts), cam_pos, lookat, up,
                            np.zeros((B, 1, 3), np.float32), np.ones((B, 1, 3), np.float32), 320, 240,
                            fov_y=g['render_fov'][:1], near_clip=np.float32(rn.CFG.fl), far_clip=np.float32(100.0))
    assert (np.abs(sil[..., 3:4] - g['render_sil'][:1]) > 1e-4).mean() < 1e-4


@pytest.mark.skipif(not rr.available(), reason='reference tree not present (GPU box): fixtures only')
def test_live_reference_reproduces_fixtures(smpl_g, tmp_path):
    """Re-runs the reference files (build container only) and checks the committed fixture is what they produce."""
    import pickle
    pkl = str(tmp_path / 'm.pkl')
    with open(pkl, 'wb') as f:
        pickle.dump(syn.make_synthetic_smpl(0), f)
    ns = rr.load(float64=True)
    tf = ns.tf
    g = smpl_g
    smpl = ns.batch_smpl.SMPL(pkl)
    sl = slice(6, 10)        # two synthetic + two SURREAL frames
    verts, joints, Rs = smpl(tf.constant(g['beta'][sl].astype(np.float64)), tf.constant(g['theta'][sl].astype(np.float64)),
                             get_skin=True)
    assert rel(verts.numpy()[:, g['vsub']], g['verts_sub'][sl]) < 1e-14
    assert rel(joints.numpy(), g['joints'][sl]) < 1e-14
    assert rel(smpl.J_transformed.numpy(), g['J_transformed'][sl]) < 1e-14
    assert 'tensorflow' not in sys.modules      # the stand-in never stays importable as `tensorflow`
