"""Generates tests/golden/reference_*.npz by RUNNING THE REFERENCE'S OWN PYTHON FILES (imported from /root/reference,
unmodified) on top of oracle/tfshim, the torch-CPU stand-in for the `tensorflow` module (TensorFlow itself is not
installable here).  These fixtures are what pins the oracle restatements for the TF-graph part of the path:

  reference_smpl.npz    tf_smpl/batch_smpl.py SMPL.__call__ (+ batch_lbs.py) and tf_perspective_projection/project.py
                        on 8 synthetic and 16 real SURREAL (pose, shape) frames, float64 arithmetic; tf.gradients of a
                        seeded linear functional of (verts, J_transformed, joints) w.r.t. (beta, theta); the same
                        forward in the reference's native float32
  reference_losses.npz  pose_3d/pose_model_3d.py PoseModel3d.get_encoder_losses (:142-334) called on a stand-in `self`:
                        every loss term (as passed to tf.summary.scalar), the total, and tf.gradients of the terms
                        w.r.t. the network output vector [B,79]
  reference_utils.npz   pose_3d/utils.py: add_axis_angle_rotations, rotate_global_pose, get_2d_points_in_3d,
                        get_camera_normal_plane, vertex_faces_from_face_verts, normals_from_mesh (real SMPL faces,
                        tf_smpl/smpl_faces.npy), render_mesh_verts_cam (the reference's renderer graph around its C++
                        rasteriser built in oracle/_ref); tf_smpl/projection.py batch_orth_proj_idrot
  reference_surreal.npz pose_3d/data_helpers.py read_maps_poses_images_surreal (:43-89) on the reference's SURREAL
                        annotation sample, synthetic heat maps / frames (pure NumPy code, no TF op involved)

The SMPL model is the synthetic seed-0 model of smplb200/synthetic.py (the licensed SMPL pickle is not available);
it is written to a temporary pickle and loaded by the reference's own SMPL.__init__.

Run from the repo root (needs /root/reference):  python tests/golden/make_reference_golden.py
"""
import os
import pickle
import sys
import tempfile
import types

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, '3d-pose-estimation_b200'))

from oracle import run_reference as rr  # noqa: E402
from smplb200 import synthetic as syn  # noqa: E402

OUT = os.path.join(ROOT, 'tests', 'golden')
VSUB = np.arange(0, 6890, 13)


def cotangents(B, seed):
    """the seeded linear functional used for the gradient fixtures (tests rebuild it from the seed)"""
    r = np.random.default_rng(seed)
    return (r.standard_normal((B, 6890, 3)) / 6890.0, r.standard_normal((B, 24, 3)), r.standard_normal((B, 19, 3)))


def smpl_inputs():
    syn_in = syn.make_inputs(8, seed=1000, dtype=np.float32)
    beta, theta = syn_in['beta'], syn_in['theta']
    cam = syn.make_inputs(24, seed=7, dtype=np.float32)
    mat = os.path.join(rr.REF, 'data', 'images', '01_11_c0001_info.mat')
    import scipy.io
    d = scipy.io.loadmat(mat)
    idx = np.arange(0, 96, 6)
    theta = np.concatenate([theta, d['pose'].T[idx].astype(np.float32)], 0)
    beta = np.concatenate([beta, d['shape'].T[idx].astype(np.float32)], 0)
    return beta, theta, cam['cam_pos'], cam['cam_rot'], cam['focal']


def run_smpl(ns, pkl, beta, theta, cam_pos, cam_rot, focal, grads):
    tf = ns.tf
    smpl = ns.batch_smpl.SMPL(pkl)
    B = beta.shape[0]
    tb, tt = tf.Variable(beta), tf.Variable(theta)
    verts, joints, Rs = smpl(tb, tt, get_skin=True)
    Jt = smpl.J_transformed
    n = ns.config.n_joints_smpl
    tile = lambda x: tf.reshape(tf.tile(x, [1, n]), [-1, 3])                     # pose_model_3d.py:202-210
    f_tile = tf.reshape(tf.tile(tf.constant(focal)[:, None], [1, n]), [-1])      # :213-215
    uv = ns.project.project(tf.reshape(Jt, [-1, 3]), tile(tf.constant(cam_pos)), tile(tf.constant(cam_rot)), f_tile)
    img_dim = tf.constant([320.0, 240.0], dtype=tf.float32)
    px = uv / ns.config.ss * img_dim[1] + img_dim / 2                           # :229
    out = dict(verts_sub=verts.numpy()[:, VSUB], verts_sum=verts.numpy().sum(1), joints=joints.numpy(),
               J_transformed=Jt.numpy(), Rs=Rs.numpy(), uv=uv.numpy().reshape(B, n, 2), px=px.numpy().reshape(B, n, 2))
    if grads:
        gv, gj, gk = cotangents(B, 4321)
        L = (tf.reduce_sum(verts * tf.constant(gv)) + tf.reduce_sum(Jt * tf.constant(gj)) +
             tf.reduce_sum(joints * tf.constant(gk)))
        g_beta, g_theta = tf.gradients(L, [tb, tt])
        out.update(g_beta=g_beta.numpy(), g_theta=g_theta.numpy(), L=float(L))
        # gradient of the pixel coordinates alone (the re-projection path): cotangent = a seeded [B,24,2]
        gp = np.random.default_rng(99).standard_normal((B, n, 2))
        Lp = tf.reduce_sum(tf.reshape(px, [B, n, 2]) * tf.constant(gp))
        gpb, gpt = tf.gradients(Lp, [tb, tt])
        out.update(g_px=gp, gpx_beta=gpb.numpy(), gpx_theta=gpt.numpy())
    return out


def make_smpl(pkl):
    beta, theta, cam_pos, cam_rot, focal = smpl_inputs()
    ns = rr.load(float64=True)
    f64 = [a.astype(np.float64) for a in (beta, theta, cam_pos, cam_rot, focal)]
    g = run_smpl(ns, pkl, *f64, grads=True)
    ns = rr.load(float64=False)
    g32 = run_smpl(ns, pkl, beta, theta, cam_pos, cam_rot, focal, grads=False)
    assert g32['verts_sub'].dtype == np.float32
    g.update({'f32_' + k: g32[k] for k in ('verts_sub', 'joints', 'J_transformed', 'px')})
    g.update(beta=beta, theta=theta, cam_pos=cam_pos, cam_rot=cam_rot, focal=focal, vsub=VSUB)
    np.savez_compressed(os.path.join(OUT, 'reference_smpl.npz'), **g)
    print('reference_smpl.npz', {k: v.shape for k, v in g.items() if hasattr(v, 'shape')})


def loss_case(ns, pkl, seed, noise, cam_noise):
    """One call of PoseModel3d.get_encoder_losses on a stand-in `self` (pose_model_3d.py:142-334)."""
    tf = ns.tf
    B = 4
    r = np.random.default_rng(seed)
    inp = syn.make_inputs(B, seed=seed, dtype=np.float32)
    gt_pose = inp['theta'].astype(np.float64)
    betas = inp['beta'].astype(np.float64)
    out_pose = gt_pose + noise * r.standard_normal(gt_pose.shape)
    cam_pos, cam_rot = inp['cam_pos'].astype(np.float64), inp['cam_rot'].astype(np.float64)
    smpl = ns.batch_smpl.SMPL(pkl)
    # ground-truth 2D joints in IMAGE coordinates (y down): projection of the GT joints by the true camera + pixel noise
    smpl(tf.constant(betas), tf.constant(gt_pose), get_skin=True)
    n = ns.config.n_joints_smpl
    tile = lambda x: tf.reshape(tf.tile(tf.constant(x), [1, n]), [-1, 3])
    uv = ns.project.project(tf.reshape(smpl.J_transformed, [-1, 3]), tile(cam_pos), tile(cam_rot),
                            tf.constant(np.full(B * n, ns.config.fl)))
    px = (uv.numpy() / ns.config.ss * 240.0 + np.array([160.0, 120.0])).reshape(B, n, 2)
    gt2d = px + r.standard_normal(px.shape)
    gt2d = (gt2d - np.array([0.0, 240.0])) * np.array([1.0, -1.0])           # undo of :196-197
    outputs = np.zeros((B, 79))
    outputs[:, :72] = out_pose
    outputs[:, 72:75] = cam_pos + cam_noise * r.standard_normal((B, 3))
    outputs[:, 75:78] = cam_rot + cam_noise * r.standard_normal((B, 3))
    outputs[:, 78] = r.standard_normal(B) * 0.1                               # raw focal output (only :290-296 reads it)
    tf.SUMMARIES.clear()
    me = types.SimpleNamespace(
        graph=tf.get_default_graph(), pose_loss=True, mesh_loss=True, reproject_loss=True, smpl=smpl,
        img_dim=tf.constant([320.0, 240.0], dtype=tf.float32), outputs=tf.Variable(outputs),
        mesh_faces=tf.constant(ns.faces, dtype=tf.int32),
        vert_faces=tf.constant(ns.utils.vertex_faces_from_face_verts(ns.faces)))
    total = ns.pose_model_3d.PoseModel3d.get_encoder_losses(
        me, me.outputs[:, :72], tf.constant(gt_pose), tf.constant(betas), tf.constant(gt2d))
    res = dict(gt_pose=gt_pose, betas=betas, gt_joints2d=gt2d, outputs=outputs, total=float(total))
    res['g_total'] = tf.gradients(total, me.outputs)[0].numpy()
    for k in ('pose_loss_direct', 'pose_loss', 'reg_loss', '3d_loss', 'reproj_loss', 'cam_loss', 'cam_angle',
              'cam_pos_diff', 'rot_dot'):
        v = tf.SUMMARIES[k]
        res['term_' + k] = float(v)
        g = tf.gradients(v, me.outputs)[0]
        res['g_' + k] = np.zeros_like(outputs) if g is None else g.numpy()
    # rays whose dot product lies between float32(1 - 1e-6) and the double 1 - 1e-6 are clipped by the float32 graph
    # but not by this float64 evaluation (:267): report them so that main() can pick a seed without such a ray
    v3 = tf.nn.l2_normalize(smpl.J_transformed - me.outputs[:, 72:75][:, tf.newaxis, :], axis=2)
    flipped = tf.constant(gt2d) * tf.constant([1.0, -1.0]) + tf.stack([0., me.img_dim[1]])
    v2 = tf.nn.l2_normalize(ns.utils.get_2d_points_in_3d(flipped, me.outputs[:, 75:78],
                                                          tf.constant(np.full(B, ns.config.fl)), me.img_dim), axis=2)
    dot = np.abs(tf.reduce_sum(v3 * v2, axis=2).numpy())
    res['n_clipped'] = int((dot > 1 - 1e-6).sum())
    res['n_ambiguous'] = int(((dot > float(np.float32(1 - 1e-6))) & (dot <= 1 - 1e-6)).sum())
    res['camera_view'] = tf.SUMMARIES['camera_view'].numpy().astype(np.float32)
    res['out_mesh'] = tf.SUMMARIES['out_mesh'].numpy().astype(np.float32)
    return res


def make_losses(pkl):
    ns = rr.load(float64=True)
    g = {}
    for name, seed, noise, cam_noise in (('near', 11, 0.02, 0.002), ('far', 12, 0.3, 0.05)):
        while True:
            c = loss_case(ns, pkl, seed, noise, cam_noise)
            if c['n_ambiguous'] == 0 and (name == 'far' or c['n_clipped'] > 0):
                break
            seed += 100
        print(name, 'seed', seed, 'clipped rays', c['n_clipped'])
        print(name, {k: round(v, 6) for k, v in c.items() if isinstance(v, float)},
              'coverage', float((c['camera_view'] > 0).mean()), float((c['out_mesh'] > 0).mean()))
        g.update({name + '_' + k: v for k, v in c.items()})
    np.savez_compressed(os.path.join(OUT, 'reference_losses.npz'), **g)


def make_utils(pkl):
    ns = rr.load(float64=True)
    tf, U = ns.tf, ns.utils
    r = np.random.default_rng(21)
    g = {}
    rv1, rv2 = r.standard_normal((64, 3)), r.standard_normal((64, 3)) * 0.7
    rv1[5] = 0.0
    rv2[9] = 0.0
    g.update(rv1=rv1, rv2=rv2, add_axis_angle=U.add_axis_angle_rotations(tf.constant(rv1), tf.constant(rv2)).numpy())
    import scipy.io
    d = scipy.io.loadmat(os.path.join(rr.REF, 'data', 'images', '01_11_c0001_info.mat'))
    thetas = d['pose'].T[::6][:16].astype(np.float64)
    zrot = r.uniform(-np.pi, np.pi, 16)
    g.update(rgp_thetas=thetas, rgp_zrot=zrot,
             rotate_global_pose=U.rotate_global_pose(tf.constant(thetas), tf.constant(zrot)).numpy())
    B = 5
    p2d, cr, cf = r.uniform(0, 320, (B, 24, 2)), r.standard_normal((B, 3)) * 0.3, r.uniform(0.03, 0.08, B)
    cp = r.standard_normal((B, 3))
    img_dim = tf.constant([320.0, 240.0], dtype=tf.float32)
    g.update(p2d=p2d, p2d_cam_rot=cr, p2d_cam_f=cf, p2d_cam_pos=cp,
             points_2d_in_3d=U.get_2d_points_in_3d(tf.constant(p2d), tf.constant(cr), tf.constant(cf), img_dim).numpy())
    nrm, dd = U.get_camera_normal_plane(tf.constant(cp), tf.constant(cr))
    g.update(cam_plane_n=nrm.numpy(), cam_plane_d=dd.numpy())
    X, cam = r.standard_normal((B, 19, 3)), r.standard_normal((B, 3))
    g.update(orth_X=X, orth_cam=cam,
             orth_proj=ns.projection.batch_orth_proj_idrot(tf.constant(X), tf.constant(cam)).numpy())
    th = r.standard_normal((3, 72)) * 0.3
    g.update(lrot_theta=th, lrotmin=ns.batch_lbs.batch_lrotmin(tf.constant(th)).numpy())
    Rs = ns.batch_lbs.batch_rodrigues(tf.constant(th.reshape(-1, 3))).numpy().reshape(3, 24, 3, 3)
    Js = r.standard_normal((3, 24, 3)) * 0.3
    parents = syn.make_synthetic_smpl(0)['kintree_table'][0].astype(np.int32)
    for rb in (False, True):
        nj, A = ns.batch_lbs.batch_global_rigid_transformation(tf.constant(Rs), tf.constant(Js), parents, rotate_base=rb)
        g.update({'grt_newJ_%d' % rb: nj.numpy(), 'grt_A_%d' % rb: A.numpy()})
    g.update(grt_Rs=Rs, grt_Js=Js)
    # mesh side: real SMPL topology, vertices of the synthetic model posed by the reference SMPL
    faces = ns.faces
    vf = U.vertex_faces_from_face_verts(faces)
    smpl = ns.batch_smpl.SMPL(pkl)
    inp = syn.make_inputs(2, seed=33, dtype=np.float32)
    verts, _, _ = smpl(tf.constant(inp['beta'].astype(np.float64)), tf.constant(inp['theta'].astype(np.float64)),
                       get_skin=True)
    normals = U.normals_from_mesh(verts, tf.constant(faces, dtype=tf.int32), tf.constant(vf))
    g.update(faces=faces.astype(np.int16), vertex_faces=vf.astype(np.int16), mesh_beta=inp['beta'], mesh_theta=inp['theta'],
             mesh_verts=verts.numpy().astype(np.float32), normals=normals.numpy())
    # the two render calls of pose_model_3d.py:316-327, in float32 like the reference
    ns = rr.load(float64=False)
    tf, U = ns.tf, ns.utils
    v32 = tf.constant(g['mesh_verts'])
    cam_pos = np.array([[0.1, -0.3, -4.0], [-0.2, -0.35, -3.5]], dtype=np.float32)
    cam_rot = np.array([[0.02, 0.05, 0.0], [0.0, -0.04, 0.01]], dtype=np.float32)
    fov = tf.atan2(ns.config.ss / 2, tf.constant(np.full(2, ns.config.fl, dtype=np.float32))) * 360 / np.pi
    tfaces, tvf = tf.constant(faces, dtype=tf.int32), tf.constant(vf)
    lit = U.render_mesh_verts_cam(v32, tf.constant(cam_pos), tf.constant(cam_rot), fov, tfaces, tvf,
                                  tf.constant(cam_pos)[:, tf.newaxis, :])
    sil = U.render_mesh_verts_cam(v32, tf.constant(cam_pos), tf.constant(cam_rot), fov, tfaces)
    fixed = U.render_mesh_verts_cam(v32, tf.constant([0.0, -0.35, -4.0]), tf.constant([0.0, 0.0, 0.0]), 30.0, tfaces,
                                    tvf, tf.constant([0.0, 1.0, -4.0]))
    g.update(render_cam_pos=cam_pos, render_cam_rot=cam_rot, render_fov=fov.numpy(), render_lit=lit.numpy(),
             render_sil=sil.numpy(), render_fixed=fixed.numpy())
    print('render coverage', float((g['render_sil'] > 0).mean()), float((g['render_lit'] > 0).mean()),
          float((g['render_fixed'] > 0).mean()))
    np.savez_compressed(os.path.join(OUT, 'reference_utils.npz'), **g)


def make_surreal(pkl):
    """pose_3d/data_helpers.py:43-89 read_maps_poses_images_surreal on the reference's own annotation sample
    (data/images/01_11_c0001_info.mat) with a synthetic heat-map file (random mask / diffs incl. NaN) and synthetic
    JPEG frames.  OpenCV is not installed: the three cv2 calls of that function are served by PIL / NumPy."""
    from PIL import Image
    cv2 = types.ModuleType('cv2')
    cv2.COLOR_BGR2RGB, cv2.NORM_MINMAX = 4, 32
    cv2.imread = lambda path: np.array(Image.open(path).convert('RGB'))[:, :, ::-1]
    cv2.cvtColor = lambda img, code: img[:, :, ::-1]
    cv2.normalize = lambda f, dst, a, b, mode: (f - f.min()) / max(float(f.max() - f.min()), 1e-12) * (b - a) + a
    ns = rr.load(float64=False, cv2=cv2)
    import scipy.io
    info = os.path.join(rr.REF, 'data', 'images', '01_11_c0001_info.mat')
    T = scipy.io.loadmat(info)['pose'].shape[1]
    r = np.random.default_rng(77)
    H, W = 6, 8        # small: the whole concat output is stored
    mask = (r.uniform(size=(T, 1)) < 0.8).astype(np.uint8)
    diffs = r.uniform(0, 300, size=(T, 1))
    diffs[r.integers(0, T, 7)] = np.nan
    tmp = tempfile.mkdtemp()
    maps = os.path.join(tmp, 'maps.mat')
    scipy.io.savemat(maps, dict(heat_mat=r.uniform(size=(H, W, 19, T)).astype(np.float32), mask=mask, diffs=diffs))
    frames = os.path.join(tmp, 'frames')
    os.mkdir(frames)
    for t in range(T):
        Image.fromarray(r.integers(0, 255, (H, W, 3), dtype=np.uint8)).save(os.path.join(frames, 'f%04d.jpg' % t))
    jpeg = [open(os.path.join(frames, 'f%04d.jpg' % t), 'rb').read() for t in range(T)]
    heat_mat = scipy.io.loadmat(maps)['heat_mat']
    with np.errstate(invalid='ignore'):
        concat, poses, shapes, joints2d, zrot = ns.data_helpers.read_maps_poses_images_surreal(
            maps, info, frames.encode('utf-8'))
    print('surreal', concat.shape, poses.shape, joints2d.dtype, zrot.shape)
    np.savez_compressed(os.path.join(OUT, 'reference_surreal.npz'), mask=mask, diffs=diffs, img_size_x=W,
                        poses=poses, shapes=shapes, joints2d=joints2d, zrot=zrot, concat=concat, heat_mat=heat_mat,
                        jpeg_blob=np.frombuffer(b''.join(jpeg), dtype=np.uint8),
                        jpeg_len=np.array([len(j) for j in jpeg], dtype=np.int64))


def make_h36m(pkl):
    """pose_3d/data_helpers.py:92-131: the reference's read_tfrecord_h36m / read_maps_h36m / concat_img_heatmaps_h36m run
    on a synthetic HMR-style TFRecord.  The Examples are serialised by the official protobuf runtime (dynamic messages
    on the public tf.train.Example schema, oracle/tfshim), which also serves tf.parse_single_example for the reference
    code; the record framing (length + masked CRC-32C) is written by smplb200.h36m.write_tfrecord."""
    import io
    from PIL import Image
    cv2 = types.ModuleType('cv2')
    ns = rr.load(float64=False, cv2=cv2)
    tf = ns.tf
    Example, _, _ = tf.example_proto_classes()
    from smplb200 import h36m
    r = np.random.default_rng(36)
    N, H, W = 6, 10, 12
    recs, fields = [], []
    for i in range(N):
        buf = io.BytesIO()
        Image.fromarray(r.integers(0, 255, (H, W, 3), dtype=np.uint8)).save(buf, format='JPEG')
        f = {'image/center': r.integers(0, 300, 2), 'image/encoded': [buf.getvalue()],
             'image/filename': [b'S1_Directions_%d.jpg' % i], 'image/format': [b'JPEG'], 'image/height': [H],
             'image/visibility': r.integers(0, 2, 14), 'image/width': [W],
             'image/x': r.uniform(0, W, 14).astype(np.float32), 'image/y': r.uniform(0, H, 14).astype(np.float32),
             'meta/crop_pt': r.integers(-50, 50, 2), 'meta/has_3d': [1],
             'meta/scale_factors': r.uniform(0.5, 2, 2).astype(np.float32),
             'mosh/gt3d': r.normal(size=42).astype(np.float32), 'mosh/pose': r.normal(0, 0.4, 72).astype(np.float32),
             'mosh/shape': r.normal(size=10).astype(np.float32)}
        ex = Example()
        for k, v in f.items():
            ft = ex.features.feature[k]
            if isinstance(v[0], (bytes, bytearray)):
                ft.bytes_list.value.extend(v)
            elif np.asarray(v).dtype.kind == 'f':
                ft.float_list.value.extend([float(x) for x in v])
            else:
                ft.int64_list.value.extend([int(x) for x in v])
        recs.append(ex.SerializeToString())
        fields.append(f)
    tmp = tempfile.mkdtemp()
    path = os.path.join(tmp, 'h36m.tfrecord')
    h36m.write_tfrecord(path, recs)
    imgs, poses, shapes, j2d = [], [], [], []
    for rec in recs:
        img, p, s, j = ns.data_helpers.read_tfrecord_h36m(rec)          # the reference function, unmodified
        imgs.append(tf._raw(img).numpy()); poses.append(tf._raw(p).numpy()); shapes.append(tf._raw(s).numpy())
        j2d.append(tf._raw(j).numpy())
    import scipy.io
    maps = os.path.join(tmp, 'maps.mat')
    scipy.io.savemat(maps, dict(heat_mat=r.uniform(size=(H, W, 19, N)).astype(np.float32)))
    heat = ns.data_helpers.read_maps_h36m(maps)
    print('h36m', np.stack(poses).shape, np.stack(j2d).shape, heat.shape, np.stack(imgs).shape)
    np.savez_compressed(os.path.join(OUT, 'reference_h36m.npz'), tfrecord=np.frombuffer(open(path, 'rb').read(), dtype=np.uint8),
                        poses=np.stack(poses), shapes=np.stack(shapes), joints2d=np.stack(j2d), imgs=np.stack(imgs),
                        heat_mat=scipy.io.loadmat(maps)['heat_mat'], heatmaps=heat)


def main():
    if not rr.available():
        raise SystemExit('needs the reference tree at ' + rr.REF)
    pkl = os.path.join(tempfile.mkdtemp(), 'synthetic_smpl_seed0.pkl')
    with open(pkl, 'wb') as f:
        pickle.dump(syn.make_synthetic_smpl(0), f)
    which = sys.argv[1:] or ['smpl', 'losses', 'utils', 'surreal', 'h36m']
    for w in which:
        {'smpl': make_smpl, 'losses': make_losses, 'utils': make_utils, 'surreal': make_surreal, 'h36m': make_h36m}[w](pkl)


if __name__ == '__main__':
    main()
