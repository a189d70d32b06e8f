"""ORACLE — TEST INFRASTRUCTURE ONLY (not product code).

Independent torch-CPU twin of `oracle/ref_numpy.py` (same reference citations): used
(a) in float64 with autograd to obtain the gradients TF autodiff would produce
    (`pose_model_3d.py:393-400`), against which the hand-written CUDA backward is checked;
(b) in float32 with all host threads as the timed CPU baseline ("port" of the reference's TF
    CPU op sequence, including tile(W) and the [N,6890,4,4] batched matmul,
    `batch_smpl.py:120-132`) in `bench.py`.
Pinned like `oracle/ref_numpy.py` (see its header): values and autograd gradients are held to the reference's own
Python files run on oracle/tfshim (tests/test_reference_pin.py).  Never imported by the product path.
"""
import math
from types import SimpleNamespace

import numpy as np
import torch

from . import ref_numpy as rn

CFG = rn.CFG


def load_model(pkl, joint_type='cocoplus', dtype=torch.float32):
    npdt = np.float64 if dtype == torch.float64 else np.float32
    m = rn.load_model(pkl, joint_type, npdt)
    t = SimpleNamespace()
    for k in ('v_template', 'shapedirs', 'J_regressor', 'posedirs', 'weights', 'joint_regressor'):
        setattr(t, k, torch.from_numpy(np.ascontiguousarray(getattr(m, k))))
    t.parents = m.parents
    t.size = m.size
    t.dtype = dtype
    return t


def batch_skew(vec):
    """batch_lbs.py:11-35."""
    N = vec.shape[0]
    z = torch.zeros(N, dtype=vec.dtype)
    return torch.stack([z, -vec[:, 2], vec[:, 1], vec[:, 2], z, -vec[:, 0],
                        -vec[:, 1], vec[:, 0], z], dim=1).reshape(N, 3, 3)


def batch_rodrigues(theta):
    """batch_lbs.py:38-57."""
    angle = torch.linalg.norm(theta + 1e-8, dim=1, keepdim=True)
    r = (theta / angle).unsqueeze(-1)
    angle = angle.unsqueeze(-1)
    cos, sin = torch.cos(angle), torch.sin(angle)
    outer = torch.matmul(r, r.transpose(1, 2))
    eyes = torch.eye(3, dtype=theta.dtype).unsqueeze(0)
    return cos * eyes + (1 - cos) * outer + sin * batch_skew(r[:, :, 0])


def batch_global_rigid_transformation(Rs, Js, parent):
    """batch_lbs.py:84-144 (rotate_base=False)."""
    N = Rs.shape[0]
    dt = Rs.dtype
    Js = Js.unsqueeze(-1)

    def make_A(R, t):
        R_homo = torch.nn.functional.pad(R, (0, 0, 0, 1))
        t_homo = torch.cat([t, torch.ones(N, 1, 1, dtype=dt)], 1)
        return torch.cat([R_homo, t_homo], 2)

    results = [make_A(Rs[:, 0], Js[:, 0])]
    for i in range(1, parent.shape[0]):
        j_here = Js[:, i] - Js[:, parent[i]]
        results.append(torch.matmul(results[parent[i]], make_A(Rs[:, i], j_here)))
    results = torch.stack(results, dim=1)
    new_J = results[:, :, :3, 3]
    Js_w0 = torch.cat([Js, torch.zeros(N, parent.shape[0], 1, 1, dtype=dt)], 2)
    init_bone = torch.matmul(results, Js_w0)
    init_bone = torch.nn.functional.pad(init_bone, (3, 0))
    return new_J, results - init_bone


def smpl_forward(m, beta, theta, tile_W=True, want_verts=True):
    """batch_smpl.py:71-143.  want_verts=False stops after J_transformed (the joint-only losses of pose_model_3d.py:194-238
    never touch the vertices; saves the checker the blend GEMM at BASELINE batch sizes)."""
    N = beta.shape[0]
    nv = m.size[0]
    dt = beta.dtype
    v_shaped = torch.matmul(beta, m.shapedirs).reshape(-1, nv, 3) + m.v_template
    J = torch.stack([torch.matmul(v_shaped[:, :, c], m.J_regressor) for c in range(3)], dim=2)
    Rs = batch_rodrigues(theta.reshape(-1, 3)).reshape(-1, 24, 3, 3)
    if not want_verts:
        J_transformed, A = batch_global_rigid_transformation(Rs, J, m.parents)
        return {'J_transformed': J_transformed, 'Rs': Rs, 'A': A}
    pose_feature = (Rs[:, 1:] - torch.eye(3, dtype=dt)).reshape(-1, 207)
    v_posed = torch.matmul(pose_feature, m.posedirs).reshape(-1, nv, 3) + v_shaped
    J_transformed, A = batch_global_rigid_transformation(Rs, J, m.parents)
    if tile_W:
        W = m.weights.repeat(N, 1).reshape(N, -1, 24)
        T = torch.matmul(W, A.reshape(N, 24, 16)).reshape(N, -1, 4, 4)
        v_posed_homo = torch.cat([v_posed, torch.ones(N, nv, 1, dtype=dt)], 2)
        verts = torch.matmul(T, v_posed_homo.unsqueeze(-1))[:, :, :3, 0]
    else:
        T = torch.einsum('vj,njk->nvk', m.weights, A.reshape(N, 24, 16)).reshape(N, nv, 4, 4)
        verts = torch.einsum('nvrc,nvc->nvr', T[:, :, :3, :3], v_posed) + T[:, :, :3, 3]
    joints = torch.stack([torch.matmul(verts[:, :, c], m.joint_regressor) for c in range(3)], dim=2)
    return dict(verts=verts, joints=joints, Rs=Rs, J_transformed=J_transformed, A=A,
                v_posed=v_posed, J=J)


def rodrigues_batch(rvecs):
    """project.py:51-81.  At exactly-zero input TF back-propagates NaN through the dead branch;
    keep test inputs away from exact zero."""
    N = rvecs.shape[0]
    dt = rvecs.dtype
    thetas = torch.linalg.norm(rvecs, dim=1, keepdim=True)
    is_zero = thetas[:, 0] == 0.0
    u = rvecs / thetas
    zero = torch.zeros(N, dtype=dt)
    Ks = torch.stack([torch.stack([zero, -u[:, 2], u[:, 1]], dim=1),
                      torch.stack([u[:, 2], zero, -u[:, 0]], dim=1),
                      torch.stack([-u[:, 1], u[:, 0], zero], dim=1)], dim=1)
    eye = torch.eye(3, dtype=dt).unsqueeze(0).expand(N, 3, 3)
    Rs = eye + torch.sin(thetas)[..., None] * Ks + (1 - torch.cos(thetas)[..., None]) * torch.matmul(Ks, Ks)
    return torch.where(is_zero[:, None, None], eye, Rs)


def project(points, cams_pos, cams_angles, focal_lens):
    """project.py:7-48."""
    N = points.shape[0]
    dt = points.dtype
    rot = rodrigues_batch(cams_angles)
    pc = torch.matmul(rot.transpose(1, 2), (points - cams_pos).unsqueeze(-1)).squeeze(-1)
    zeros, ones = torch.zeros(N, dtype=dt), torch.ones(N, dtype=dt)
    homog = torch.cat([pc, ones[:, None]], dim=1).unsqueeze(-1)
    pm = torch.stack([torch.stack([focal_lens, zeros, zeros, zeros], dim=1),
                      torch.stack([zeros, focal_lens, zeros, zeros], dim=1),
                      torch.stack([zeros, zeros, ones, zeros], dim=1),
                      torch.stack([zeros, zeros, ones, zeros], dim=1)], dim=1)
    p2h = torch.matmul(pm, homog).squeeze(-1)
    return (p2h / p2h[:, 3, None])[:, :2]


def tile_cam(x, n=24):
    if x.dim() == 1:
        return x[:, None].repeat(1, n).reshape(-1)
    return x.repeat(1, n).reshape(-1, x.shape[1])


def project_joints_px(joints, cam_pos, cam_rot, focal, img_w=320.0, img_h=240.0, ss=CFG.ss):
    """pose_model_3d.py:202-231."""
    B, P = joints.shape[:2]
    uv = project(joints.reshape(-1, 3), tile_cam(cam_pos, P), tile_cam(cam_rot, P), tile_cam(focal, P))
    img_dim = torch.tensor([img_w, img_h], dtype=uv.dtype)
    px = uv / ss * img_dim[1] + img_dim / 2
    return uv.reshape(B, P, 2), px.reshape(B, P, 2)


def mse(labels, predictions, weight=1.0):
    return weight * torch.mean((predictions - labels) ** 2)


def reproj_loss(out_joints, gt_joints3d, cam_pos, cam_rot, focal, img_w=320.0, img_h=240.0,
                ss=CFG.ss, scale=None, cap=True):
    """pose_model_3d.py:218-238; camera detached (stop_gradient :218-221)."""
    if scale is None:
        scale = 1.0 * ss
    cp, cr, f = cam_pos.detach(), cam_rot.detach(), focal.detach()
    _, out_2d = project_joints_px(out_joints, cp, cr, f, img_w, img_h, ss)
    _, gt_2d = project_joints_px(gt_joints3d, cp, cr, f, img_w, img_h, ss)
    loss = mse(gt_2d, out_2d, scale)
    if cap:
        loss = torch.minimum(torch.tensor(1.0 / CFG.n_joints_smpl, dtype=loss.dtype), loss)
    return loss, out_2d, gt_2d


def cam_loss(gt_joints3d, gt_joints2d_flipped, cam_pos, cam_rot, focal, img_w=320.0, img_h=240.0,
             ss=CFG.ss, scale=CFG.cam_loss_scale, cap=True):
    """pose_model_3d.py:240-254."""
    _, px = project_joints_px(gt_joints3d, cam_pos, cam_rot, focal, img_w, img_h, ss)
    loss = mse(gt_joints2d_flipped, px, scale)
    if cap:
        loss = torch.minimum(torch.tensor(0.5, dtype=loss.dtype), loss)
    return loss, px


def loss3d(gt_meshes, out_meshes, gt_joints3d, out_joints):
    """pose_model_3d.py:183-192."""
    return mse(gt_meshes, out_meshes, CFG.mesh_loss_scale) + mse(gt_joints3d, out_joints, CFG.joint_loss_scale)


# ---- N3: camera-angle loss and regularisers (pose_model_3d.py:256-294), autograd twin ----

def l2_normalize(x, dim, eps=1e-12):
    return x / torch.sqrt(torch.clamp((x * x).sum(dim, keepdim=True), min=eps))


def get_2d_points_in_3d(points_2d, cam_rot, cam_f, img_w=320.0, img_h=240.0, ss=CFG.ss):
    img_dim = torch.tensor([img_w, img_h], dtype=points_2d.dtype)
    rescaled = ss * (points_2d - img_dim / 2) / img_dim[1]
    fl = cam_f.reshape(-1, 1, 1).expand(-1, points_2d.shape[1], 1)
    p = torch.cat([rescaled, fl], dim=2)
    R = rodrigues_batch(cam_rot)[:, None]
    return torch.matmul(R, p[..., None])[..., 0]


def cam_angle_loss(gt_joints3d, gt_joints2d, cam_pos, cam_rot, cam_f, img_w=320.0, img_h=240.0, ss=CFG.ss,
                   scale=CFG.cam_angle_loss_scale):
    v3 = l2_normalize(gt_joints3d - cam_pos[:, None, :], 2)
    v2 = l2_normalize(get_2d_points_in_3d(gt_joints2d, cam_rot, cam_f, img_w, img_h, ss), 2)
    dot = torch.clamp((v3 * v2).sum(2), -1.0 + 1e-6, 1.0 - 1e-6)
    ang = torch.acos(dot)
    return scale * (ang * ang).mean()


def cam_reg_losses(cam_pos, cam_rot, f_raw):
    lim = torch.stack([-cam_pos[:, 2] - 0.3, cam_pos[:, 2] + 10.0, f_raw], dim=1)
    neg = torch.where(lim < 0, lim, torch.zeros_like(lim))
    nrm = torch.linalg.norm(cam_rot.reshape(-1, 3), dim=1)
    big = torch.where(nrm > math.pi * 0.125, nrm, torch.zeros_like(nrm))
    return (neg * neg).mean(), 10.0 * (big * big).mean()


def _l2n(x, dim):
    """tf.nn.l2_normalize: x * rsqrt(max(sum(x^2), 1e-12))."""
    return x / torch.sqrt(torch.clamp((x * x).sum(dim, keepdim=True), min=1e-12))


def phong_shader(normals, alphas, pixel_positions, light_positions, light_intensities, diffuse_colors=None,
                 camera_position=None, specular_colors=None, shininess_coefficients=None, ambient_color=None):
    """deps/tf_mesh_renderer/mesh_renderer.py:23-162 on torch tensors (any dtype; autograd gives the gradients TF's
    would): [B,H,W,*] buffers -> lit RGBA [B,H,W,4], rows reversed.  Checker of csrc/shade.cu."""
    B, H, W = normals.shape[:3]
    L = light_positions.shape[1]
    normals = normals.reshape(B, -1, 3)
    alphas = alphas.reshape(B, -1, 1)
    diffuse_colors = diffuse_colors.reshape(B, -1, 3)
    out = torch.zeros((B, H * W, 3), dtype=normals.dtype, device=normals.device)
    if ambient_color is not None:
        out = out + ambient_color.unsqueeze(1) * diffuse_colors                                # :92-94
    pos = pixel_positions.reshape(B, -1, 3)
    to_l = _l2n(light_positions.unsqueeze(2) - pos.unsqueeze(1), 3)                            # :100-102
    ndl = torch.clamp((normals.unsqueeze(1) * to_l).sum(3), 0.0, 1.0)                          # :106-109
    out = out + (diffuse_colors.unsqueeze(1) * ndl.unsqueeze(3) * light_intensities.unsqueeze(2)).sum(1)   # :112-118
    if camera_position is not None:
        spec = specular_colors.reshape(B, -1, 3)
        mirror = _l2n(2.0 * ndl.unsqueeze(3) * normals.unsqueeze(1) - to_l, 3)                 # :123-126
        to_cam = _l2n(camera_position.reshape(B, 1, 3) - pos, 2)                               # :127-128
        rdc = (to_cam.unsqueeze(1) * mirror).sum(3)                                            # :129-132
        rdc = torch.clamp(_l2n(rdc, 2), 0.0, 1.0)                                              # :134-136 (over pixels!)
        rdc = torch.where(ndl != 0.0, rdc, torch.zeros_like(rdc)).reshape(B, L, H, W)          # :140-146
        shin = torch.as_tensor(shininess_coefficients, dtype=normals.dtype, device=normals.device)
        shin = shin if shin.dim() == 3 else shin.reshape(-1, 1, 1)
        specularity = torch.pow(rdc, shin.unsqueeze(1)).reshape(B, L, H * W, 1)                # :148-152
        out = out + (spec.unsqueeze(1) * specularity * light_intensities.unsqueeze(2)).sum(1)  # :153-157
    rgb = out.reshape(B, H, W, 3)
    a = alphas.reshape(B, H, W, 1)
    rgb = torch.where((a > 0.5).expand(-1, -1, -1, 3), rgb, torch.zeros_like(rgb))             # :160-161
    return torch.flip(torch.cat([rgb, a], dim=3), dims=[1])                                    # :162


def clip_transform(vertices, projection_matrices):
    """deps/tf_mesh_renderer/rasterize_triangles.py:73-86 on torch tensors: -> (ndc [B,V,3], guarded clip w [B,V])."""
    B, V = vertices.shape[:2]
    vh = torch.cat([vertices, torch.ones((B, V, 1), dtype=vertices.dtype, device=vertices.device)], 2)   # :73-76
    clip = torch.matmul(vh, projection_matrices.transpose(1, 2))                                          # :77-79
    w = clip[:, :, 3:4]
    clip_w = torch.clamp(w.abs(), min=1e-6) * torch.sign(w)                                               # :81-84
    return clip[:, :, 0:3] / clip_w, clip_w[:, :, 0]                                                      # :85-86


def interpolate(attributes, clip_w, triangles, bary, ids, background_value):
    """rasterize_triangles.py:119-164 on torch tensors: perspective-correct barycentrics, attribute interpolation, alpha
    from the barycentric sum, background compositing.  triangles [T,3] long, bary [B,H,W,3], ids [B,H,W] long."""
    B, V = attributes.shape[:2]
    H, W = ids.shape[1:]
    dev = attributes.device
    vids = triangles[ids.reshape(B, -1)] + (torch.arange(B, device=dev) * V).view(B, 1, 1)               # :119-124
    vids = vids.reshape(-1, 3)
    b = bary.reshape(-1, 3)
    corner_attr = attributes.reshape(B * V, -1)[vids]
    corner_w = torch.reciprocal(clip_w.reshape(-1))[vids]                                                # :125-129
    weighted = b * corner_w                                                                              # :133-135
    factor = weighted.sum(1)
    corrected = weighted / torch.clamp(factor, min=1e-6).unsqueeze(1)                                    # :139-144
    images = (corner_attr * corrected.unsqueeze(2)).sum(1).reshape(B, H, W, -1)                          # :147-152
    alphas = torch.clamp((2.0 * corrected).sum(1), 0.0, 1.0).reshape(B, H, W, 1)                         # :157-159
    return alphas * images + (1.0 - alphas) * background_value                                           # :161-164
