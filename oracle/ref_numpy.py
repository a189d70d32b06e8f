"""ORACLE — TEST INFRASTRUCTURE ONLY (not product code).

CPU NumPy restatement of the reference's batched SMPL forward, perspective
projection and 2D re-projection loss.  Only `tests/`, `__graft_entry__.smoke()`
and `bench.py`'s cpu_baseline / `--impl reference` leg may import this module,
and only as the checker (or the timed CPU baseline) — never the product path.

PARITY PIN: the reference (blzq/3D-pose-estimation) is TensorFlow-1 graph code with no test, golden vector or
fixture for this path, and TensorFlow is not installable in this image.  The restatement is
pinned on the REFERENCE'S OWN PYTHON FILES instead: `oracle/run_reference.py` imports them unmodified from
/root/reference and executes them on `oracle/tfshim` (a torch-CPU stand-in for the `tensorflow` module -- TF's op
library is substituted, the reference's statement sequence is not); `tests/golden/make_reference_golden.py` stores
their float64 outputs and tf.gradients, and `tests/test_reference_pin.py` holds every function below to them at
1e-12 (forward), 1e-11 (N2 helpers), 1e-8 (gradients of the torch twin).  What stays unpinned is TensorFlow's own
float32 kernel rounding (summation order inside tf.matmul), which the tolerance of the GPU tests absorbs.
Further anchors: (i) the hand-derived values of the one example in `project.py:92-97`, (ii) analytic identities,
(iii) the independent torch float64 twin with autograd (`oracle/ref_torch.py`) and finite differences.

All functions are dtype-generic: pass float32 arrays for the "fp32 restatement"
(the reference runs fp32, `batch_smpl.py:17`) or float64 for the high-precision
oracle the CUDA path is compared against.
"""
import pickle
from types import SimpleNamespace

import numpy as np

# deps/pose_3d/config.py:6-24
CFG = SimpleNamespace(
    n_joints_smpl=24, input_img_size=(240, 320), fl=0.05, ss=0.024,
    total_loss_scale=1.0, mesh_loss_scale=1.0, joint_loss_scale=2.0,
    reproj_loss_scale=1.0 * 0.024, cam_loss_scale=1.0 / (400 * 24), cam_angle_loss_scale=10.0)


def load_model(pkl, joint_type='cocoplus', dtype=np.float32):
    """deps/tf_smpl/batch_smpl.py:17-69 (SMPL.__init__)."""
    if isinstance(pkl, dict):
        dd = pkl
    else:
        with open(pkl, 'rb') as f:
            dd = pickle.load(f, encoding='latin1')
    m = SimpleNamespace()
    m.v_template = np.asarray(dd['v_template'], dtype=dtype)               # :25-28
    m.size = [m.v_template.shape[0], 3]                                    # :30
    m.num_betas = dd['shapedirs'].shape[-1]                                # :31
    m.shapedirs = np.reshape(dd['shapedirs'], [-1, m.num_betas]).T.astype(dtype)   # :34-37 [10,20670]
    m.J_regressor = np.asarray(dd['J_regressor'].T.todense(), dtype=dtype)  # :40-43 [6890,24]
    num_pose_basis = dd['posedirs'].shape[-1]                              # :46
    m.posedirs = np.reshape(dd['posedirs'], [-1, num_pose_basis]).T.astype(dtype)  # :48-51 [207,20670]
    m.parents = dd['kintree_table'][0].astype(np.int32)                    # :54
    m.weights = np.asarray(dd['weights'], dtype=dtype)                     # :57-60
    m.joint_regressor = np.asarray(dd['cocoplus_regressor'].T.todense(), dtype=dtype)  # :63-66
    if joint_type == 'lsp':                                                # :68-69
        m.joint_regressor = m.joint_regressor[:, :14]
    m.dtype = dtype
    return m


def batch_skew(vec, batch_size=None):
    """deps/tf_smpl/batch_lbs.py:11-35 — scatter into flat 3x3 at {1,2,3,5,6,7}."""
    if batch_size is None:
        batch_size = vec.shape[0]
    col_inds = np.array([1, 2, 3, 5, 6, 7])
    indices = (np.arange(batch_size)[:, None] * 9 + col_inds).reshape(-1)
    updates = np.stack(
        [-vec[:, 2], vec[:, 1], vec[:, 2], -vec[:, 0], -vec[:, 1], vec[:, 0]], axis=1).reshape(-1)
    res = np.zeros(batch_size * 9, dtype=vec.dtype)
    res[indices] = updates
    return res.reshape(batch_size, 3, 3)


def batch_rodrigues(theta):
    """deps/tf_smpl/batch_lbs.py:38-57.  eps enters the norm only (:45), r = theta/angle (:46)."""
    dt = theta.dtype
    batch_size = theta.shape[0]
    angle = np.linalg.norm(theta + dt.type(1e-8), axis=1)[:, None]         # :45
    r = (theta / angle)[:, :, None]                                        # :46
    angle = angle[:, :, None]
    cos = np.cos(angle)
    sin = np.sin(angle)
    outer = np.matmul(r, np.transpose(r, (0, 2, 1)))                       # :52
    eyes = np.tile(np.eye(3, dtype=dt)[None], (batch_size, 1, 1))
    R = cos * eyes + (1 - cos) * outer + sin * batch_skew(r[:, :, 0], batch_size=batch_size)  # :55-56
    return R.astype(dt)


def batch_global_rigid_transformation(Rs, Js, parent, rotate_base=False):
    """deps/tf_smpl/batch_lbs.py:84-144."""
    dt = Rs.dtype
    N = Rs.shape[0]
    if rotate_base:                                                        # :102-106
        rot_x = np.array([[1, 0, 0], [0, -1, 0], [0, 0, -1]], dtype=dt)
        root_rotation = np.matmul(Rs[:, 0], np.tile(rot_x[None], (N, 1, 1)))
    else:
        root_rotation = Rs[:, 0]
    Js = Js[..., None]                                                     # :111

    def make_A(R, t):                                                      # :113-118
        R_homo = np.pad(R, [[0, 0], [0, 1], [0, 0]])
        t_homo = np.concatenate([t, np.ones([N, 1, 1], dtype=dt)], 1)
        return np.concatenate([R_homo, t_homo], 2)

    A0 = make_A(root_rotation, Js[:, 0])                                   # :120
    results = [A0]
    for i in range(1, parent.shape[0]):                                    # :122-127
        j_here = Js[:, i] - Js[:, parent[i]]
        A_here = make_A(Rs[:, i], j_here)
        results.append(np.matmul(results[parent[i]], A_here))
    results = np.stack(results, axis=1)                                    # :130
    new_J = results[:, :, :3, 3]                                           # :132
    Js_w0 = np.concatenate([Js, np.zeros([N, parent.shape[0], 1, 1], dtype=dt)], 2)   # :138
    init_bone = np.matmul(results, Js_w0)                                  # :139
    init_bone = np.pad(init_bone, [[0, 0], [0, 0], [0, 0], [3, 0]])        # :141
    A = results - init_bone                                                # :142
    return new_J, A


def smpl_forward(m, beta, theta, tile_W=True):
    """deps/tf_smpl/batch_smpl.py:71-143 (SMPL.__call__).  Returns every intermediate.

    tile_W=True keeps the reference's op sequence (tile(weights) -> [N,6890,24], batched matmul
    with A reshaped [N,24,16], then the [N,6890,4,4]x[N,6890,4,1] product, :120-132).
    tile_W=False contracts the same sums without materialising W/T (identical maths; used for
    larger batches in tests).
    """
    dt = m.dtype
    beta = np.asarray(beta, dtype=dt)
    theta = np.asarray(theta, dtype=dt)
    N = beta.shape[0]
    nv = m.size[0]
    v_shaped = np.reshape(np.matmul(beta, m.shapedirs), [-1, nv, 3]) + m.v_template      # :93-95
    Jx = np.matmul(v_shaped[:, :, 0], m.J_regressor)                                     # :98-100
    Jy = np.matmul(v_shaped[:, :, 1], m.J_regressor)
    Jz = np.matmul(v_shaped[:, :, 2], m.J_regressor)
    J = np.stack([Jx, Jy, Jz], axis=2)                                                   # :101
    Rs = np.reshape(batch_rodrigues(np.reshape(theta, [-1, 3])), [-1, 24, 3, 3])         # :105-106
    pose_feature = np.reshape(Rs[:, 1:] - np.eye(3, dtype=dt), [-1, 207])                # :109-110
    v_posed = np.reshape(np.matmul(pose_feature, m.posedirs), [-1, nv, 3]) + v_shaped    # :113-115
    J_transformed, A = batch_global_rigid_transformation(Rs, J, m.parents)               # :118
    if tile_W:
        W = np.reshape(np.tile(m.weights, [N, 1]), [N, -1, 24])                          # :122-123
        T = np.reshape(np.matmul(W, np.reshape(A, [N, 24, 16])), [N, -1, 4, 4])          # :125-127
        v_posed_homo = np.concatenate([v_posed, np.ones([N, nv, 1], dtype=dt)], 2)       # :128-129
        v_homo = np.matmul(T, v_posed_homo[..., None])                                   # :130
        verts = v_homo[:, :, :3, 0]                                                      # :132
    else:
        T = np.einsum('vj,njk->nvk', m.weights, np.reshape(A, [N, 24, 16])).reshape(N, nv, 4, 4)
        verts = np.einsum('nvrc,nvc->nvr', T[:, :, :3, :3], v_posed) + T[:, :, :3, 3]
    joint_x = np.matmul(verts[:, :, 0], m.joint_regressor)                               # :135-137
    joint_y = np.matmul(verts[:, :, 1], m.joint_regressor)
    joint_z = np.matmul(verts[:, :, 2], m.joint_regressor)
    joints = np.stack([joint_x, joint_y, joint_z], axis=2)                               # :138
    return dict(verts=verts, joints=joints, Rs=Rs, J_transformed=J_transformed,
                J=J, A=A, v_posed=v_posed, v_shaped=v_shaped, pose_feature=pose_feature)


def batch_orth_proj_idrot(X, camera):
    """deps/tf_smpl/projection.py:10-26."""
    camera = np.reshape(camera, [-1, 1, 3])
    X_trans = X[:, :, :2] + camera[:, :, 1:]
    shape = X_trans.shape
    return np.reshape(camera[:, :, 0] * np.reshape(X_trans, [shape[0], -1]), shape)


def rodrigues_batch(rvecs):
    """deps/tf_perspective_projection/project.py:51-81 (K^2 form, exact-zero guard)."""
    dt = rvecs.dtype
    batch_size = rvecs.shape[0]
    thetas = np.linalg.norm(rvecs, axis=1, keepdims=True)                  # :62
    is_zero = (thetas[:, 0] == 0.0)                                        # :63
    with np.errstate(divide='ignore', invalid='ignore'):
        u = rvecs / thetas                                                 # :64
    zero = np.zeros([batch_size], dtype=dt)
    Ks_1 = np.stack([zero, -u[:, 2], u[:, 1]], axis=1)                     # :69-71
    Ks_2 = np.stack([u[:, 2], zero, -u[:, 0]], axis=1)
    Ks_3 = np.stack([-u[:, 1], u[:, 0], zero], axis=1)
    Ks = np.stack([Ks_1, Ks_2, Ks_3], axis=1)                              # :73
    eye = np.tile(np.eye(3, dtype=dt)[None], (batch_size, 1, 1))
    Rs = eye + np.sin(thetas)[..., None] * Ks + \
        (1 - np.cos(thetas)[..., None]) * np.matmul(Ks, Ks)               # :75-77
    return np.where(is_zero[:, None, None], eye, Rs).astype(dt)            # :80-81


def project(points_world_space, cams_pos, cams_angles, focal_lens):
    """deps/tf_perspective_projection/project.py:7-48."""
    dt = points_world_space.dtype
    batch_size = points_world_space.shape[0]
    rot_mats = rodrigues_batch(cams_angles)                                # :19
    points_cam_space = np.matmul(np.transpose(rot_mats, (0, 2, 1)),
                                 (points_world_space - cams_pos)[:, :, None])[:, :, 0]   # :22-25
    zeros = np.zeros([batch_size], dtype=dt)
    ones = np.ones([batch_size], dtype=dt)
    points_cam_homog = np.concatenate([points_cam_space, ones[:, None]], axis=1)[:, :, None]  # :30-31
    proj_mat_1 = np.stack([focal_lens, zeros, zeros, zeros], axis=1)       # :35-38
    proj_mat_2 = np.stack([zeros, focal_lens, zeros, zeros], axis=1)
    proj_mat_3 = np.stack([zeros, zeros, ones, zeros], axis=1)
    proj_mat_4 = np.stack([zeros, zeros, ones, zeros], axis=1)
    proj_mat = np.stack([proj_mat_1, proj_mat_2, proj_mat_3, proj_mat_4], axis=1)   # :40-41
    points_2d_homog = np.matmul(proj_mat, points_cam_homog)[:, :, 0]       # :42-43
    points_2d = points_2d_homog / points_2d_homog[:, 3, None]              # :46
    return points_2d[:, :2]                                                # :48


def to_pixels(uv, img_w=320.0, img_h=240.0, ss=CFG.ss):
    """deps/pose_3d/pose_model_3d.py:229-231: `uv / ss * img_dim[1] + img_dim / 2`, img_dim=[W,H] (:41-42)."""
    img_dim = np.array([img_w, img_h], dtype=uv.dtype)
    return uv / uv.dtype.type(ss) * img_dim[1] + img_dim / 2


def tile_cam(x, n=24):
    """pose_model_3d.py:202-213: tile per-sample camera params over the 24 joints."""
    if x.ndim == 1:
        return np.reshape(np.tile(x[:, None], [1, n]), [-1])
    return np.reshape(np.tile(x, [1, n]), [-1, x.shape[1]])


def project_joints_px(joints, cam_pos, cam_rot, focal, img_w=320.0, img_h=240.0, ss=CFG.ss):
    """project() of [B,24,3] joints with per-sample cameras tiled per joint + pixel rescale
    (pose_model_3d.py:202-231).  Returns (uv[B,24,2], px[B,24,2])."""
    B, P = joints.shape[:2]
    uv = project(np.reshape(joints, [-1, 3]), tile_cam(cam_pos, P), tile_cam(cam_rot, P),
                 tile_cam(focal, P))
    px = to_pixels(uv, img_w, img_h, ss)
    return uv.reshape(B, P, 2), px.reshape(B, P, 2)


def mse(labels, predictions, weight=1.0):
    """tf.losses.mean_squared_error with scalar weight: weight * mean over all elements."""
    d = predictions - labels
    return d.dtype.type(weight) * np.mean(d * d)


def reproj_loss(out_joints, gt_joints3d, cam_pos, cam_rot, focal, img_w=320.0, img_h=240.0,
                ss=CFG.ss, scale=None, cap=True):
    """deps/pose_3d/pose_model_3d.py:218-238 (camera is stop_gradient'ed there, :218-221)."""
    if scale is None:
        scale = 1.0 * ss                                                   # config.py:22
    _, out_2d = project_joints_px(out_joints, cam_pos, cam_rot, focal, img_w, img_h, ss)
    _, gt_2d = project_joints_px(gt_joints3d, cam_pos, cam_rot, focal, img_w, img_h, ss)
    loss = mse(gt_2d.reshape(-1, 2), out_2d.reshape(-1, 2), scale)         # :233-235
    if cap:
        loss = np.minimum(out_2d.dtype.type(1.0 / CFG.n_joints_smpl), loss)   # :237
    return loss, out_2d, gt_2d


def flip_gt2d(gt_joints2d, img_h=240.0):
    """pose_model_3d.py:196-197: y <- H - y."""
    return gt_joints2d * np.array([1.0, -1.0], dtype=gt_joints2d.dtype) + \
        np.array([0.0, img_h], dtype=gt_joints2d.dtype)


def cam_loss(gt_joints3d, gt_joints2d_flipped, cam_pos, cam_rot, focal, img_w=320.0, img_h=240.0,
             ss=CFG.ss, scale=CFG.cam_loss_scale, cap=True):
    """deps/pose_3d/pose_model_3d.py:240-254 (gradient flows to the camera here)."""
    _, px = project_joints_px(gt_joints3d, cam_pos, cam_rot, focal, img_w, img_h, ss)
    loss = mse(gt_joints2d_flipped.reshape(-1, 2), px.reshape(-1, 2), scale)
    if cap:
        loss = np.minimum(px.dtype.type(0.5), loss)
    return loss, px


def loss3d(gt_meshes, out_meshes, gt_joints3d, out_joints):
    """deps/pose_3d/pose_model_3d.py:183-192."""
    return mse(gt_meshes, out_meshes, CFG.mesh_loss_scale) + \
        mse(gt_joints3d, out_joints, CFG.joint_loss_scale)


def adam_tf1(x, g, m, v, t, lr=2e-4, b1=0.9, b2=0.999, eps=1e-8):
    """tf.train.AdamOptimizer update (pose_model_3d.py:393): eps added to sqrt(v), lr_t carries
    both bias corrections.  t is the 1-based step count.  Returns (x, m, v)."""
    dt = x.dtype.type
    lr_t = dt(lr) * np.sqrt(dt(1) - dt(b2) ** dt(t)) / (dt(1) - dt(b1) ** dt(t))
    m = dt(b1) * m + dt(1 - b1) * g
    v = dt(b2) * v + dt(1 - b2) * g * g
    x = x - lr_t * m / (np.sqrt(v) + dt(eps))
    return x, m, v


# ---- helpers for the N2 "next" row (deps/pose_3d/utils.py:152-199) -------------------------

def add_axis_angle_rotations(rv1, rv2):
    """deps/pose_3d/utils.py:170-199."""
    a = np.linalg.norm(rv1, axis=1, keepdims=True)
    b = np.linalg.norm(rv2, axis=1, keepdims=True)
    with np.errstate(divide='ignore', invalid='ignore'):
        l = rv1 / a
        m = rv2 / b
        cos_half_c = np.cos(a / 2) * np.cos(b / 2) - (
            np.sin(a / 2) * np.sin(b / 2) * np.sum(l * m, axis=1, keepdims=True))
        sin_half_c_n = np.sin(a / 2) * np.cos(b / 2) * l + (
            np.cos(a / 2) * np.sin(b / 2) * m + np.sin(a / 2) * np.sin(b / 2) * np.cross(l, m))
        half_c = np.arccos(cos_half_c)
        n = sin_half_c_n / np.sin(half_c)
        result = n * half_c * 2
    result = np.where((a[:, 0] == 0.0)[:, None], rv2, result)
    result = np.where((b[:, 0] == 0.0)[:, None], rv1, result)
    return result.astype(rv1.dtype)


def rotate_global_pose(thetas, zrot):
    """deps/pose_3d/utils.py:152-167."""
    dt = thetas.dtype
    B = thetas.shape[0]
    turn_x_batch = np.tile(np.array([-np.pi / 2, 0.0, 0.0], dtype=dt)[None], (B, 1))
    zeros = np.zeros(B, dtype=dt)
    turn_y_batch = np.stack([zeros, zrot + dt.type(np.pi / 2), zeros], axis=-1)
    turn_both = add_axis_angle_rotations(turn_y_batch, turn_x_batch)
    g = add_axis_angle_rotations(turn_both, thetas[:, :3])
    return np.concatenate([g, thetas[:, 3:]], axis=1)


# ---- N3 "next" row: camera-angle (ray) loss and camera regularisers (deps/pose_3d/pose_model_3d.py:256-294) ----

def l2_normalize(x, axis, eps=1e-12):
    """tf.nn.l2_normalize: x * rsqrt(max(sum(x^2), eps))."""
    return x / np.sqrt(np.maximum(np.sum(x * x, axis=axis, keepdims=True), x.dtype.type(eps)))


def get_2d_points_in_3d(points_2d, cam_rot, cam_f, img_w=320.0, img_h=240.0, ss=CFG.ss):
    """deps/pose_3d/utils.py:75-89: pixels -> points on the camera image plane, rotated into world axes."""
    dt = points_2d.dtype
    img_dim = np.array([img_w, img_h], dtype=dt)
    rescaled_2d = dt.type(ss) * (points_2d - img_dim / 2) / img_dim[1]                   # :77
    fl = np.tile(np.reshape(cam_f, [-1, 1, 1]), [1, points_2d.shape[1], 1])             # :78-79
    p = np.concatenate([rescaled_2d, fl], axis=2)                                        # :80
    R = rodrigues_batch(cam_rot)[:, None]                                                # :81-83
    return np.matmul(R, p[..., None])[..., 0]                                            # :84-86


def cam_angle_loss(gt_joints3d, gt_joints2d, cam_pos, cam_rot, cam_f, img_w=320.0, img_h=240.0, ss=CFG.ss,
                   scale=CFG.cam_angle_loss_scale):
    """pose_model_3d.py:256-270."""
    dt = gt_joints3d.dtype
    v3 = l2_normalize(gt_joints3d - cam_pos[:, None, :], 2)                              # :260-261
    v2 = l2_normalize(get_2d_points_in_3d(gt_joints2d, cam_rot, cam_f, img_w, img_h, ss), 2)   # :262-264
    dot = np.sum(v3 * v2, axis=2)                                                        # :265
    dot = np.clip(dot, dt.type(-1.0 + 1e-6), dt.type(1.0 - 1e-6))                        # :267
    ang = np.arccos(dot)                                                                 # :268
    return mse(np.zeros_like(ang), ang, scale)                                           # :269-271


def cam_reg_losses(cam_pos, cam_rot, f_raw):
    """pose_model_3d.py:273-291: (cam_neg_loss, cam_rot_loss); outputs[:,74] = camera z, outputs[:,78] = raw f."""
    dt = cam_pos.dtype
    lim = np.stack([cam_pos[:, 2], cam_pos[:, 2], f_raw], axis=1)                        # :276
    lim = lim * np.array([-1.0, 1.0, 1.0], dtype=dt)[None] + np.array([-0.3, 10.0, 0.0], dtype=dt)   # :277-278
    neg = np.where(lim < 0, lim, np.zeros_like(lim))                                     # :279-281
    neg_loss = mse(np.zeros_like(neg), neg)                                              # :282-283
    nrm = np.linalg.norm(np.reshape(cam_rot, [-1, 3]), axis=1)                           # :285-286
    big = np.where(nrm > dt.type(np.pi * 0.125), nrm, np.zeros_like(nrm))                # :288-289
    rot_loss = mse(np.zeros_like(big), big, 10.0)                                        # :290-292
    return neg_loss, rot_loss


# ---- N1 (shading input): vertex normals, deps/pose_3d/utils.py:92-149 ------------------------------------------

def vertex_faces_from_face_verts(faces):
    """utils.py:130-149: adjacency lists padded with the index n_vertices."""
    n_vertices = int(np.amax(faces)) + 1
    vertex_faces = [[] for _ in range(n_vertices)]
    for face_idx, face in enumerate(faces):
        vertex_faces[face[0]].append(face_idx)
        vertex_faces[face[1]].append(face_idx)
        vertex_faces[face[2]].append(face_idx)
    orders = [len(v) for v in vertex_faces]
    mo = max(orders)
    return np.array([vf + [n_vertices] * (mo - vo) for vf, vo in zip(vertex_faces, orders)])


def normals_from_mesh(vertices, faces, vertex_faces):
    """utils.py:92-127, as written (face normals normalised over the triangle axis, :109-110; out-of-range gather
    indices give zeros as on the GPU, in-range pad indices gather that face)."""
    dt = vertices.dtype
    v1, v2, v3 = vertices[:, faces[:, 0]], vertices[:, faces[:, 1]], vertices[:, faces[:, 2]]   # :103-105
    fn = np.cross(v2 - v1, v3 - v1)                                                             # :107
    fn = l2_normalize(fn, 1)                                                                    # :108 (axis=1!)
    T = fn.shape[1]
    fn_pad = np.concatenate([fn, np.zeros((fn.shape[0], 1, 3), dtype=dt)], axis=1)
    idx = np.where(vertex_faces < T, vertex_faces, T)
    normals = fn_pad[:, idx].sum(axis=2)                                                        # :110-111
    normals = l2_normalize(normals, 2)                                                          # :112
    centered = vertices - vertices.mean(axis=1, keepdims=True)                                  # :115-116
    s = np.sum(centered * normals, axis=1)                                                      # :117
    gt, lt = np.count_nonzero(s > 0, axis=1), np.count_nonzero(s < 0, axis=1)                   # :119-120
    sign = 2 * (gt > lt).astype(dt) - 1                                                         # :122
    return normals * sign.reshape(-1, 1, 1)
