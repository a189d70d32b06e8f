"""TEST INFRASTRUCTURE ONLY (oracle): NumPy restatement of the reference's rendering graph around the rasteriser op,
statement for statement with file:line citations (relative to /root/reference/deps/tf_mesh_renderer/):
camera_utils.py (perspective :22-51, look_at :54-113, euler_matrices :116-147), rasterize_triangles.py:30-178,
mesh_renderer.py (phong_shader :23-162, tone_mapper :165-194, mesh_renderer :197-401).  The RasterizeTriangles op
itself is oracle/raster.py (the reference's C++ when built, else the C restatement).

Pinned by tests/test_render_oracle.py against the reference's golden renders (test_data/Gray_Cube_{0,1}.png,
Colored_Cube_{0,1}.png; scenes of mesh_renderer_test.py:45-166) with the reference's own comparison rule
(test_utils.py:80-121).  Only tests/, smoke() and bench.py's CPU legs may import this module."""
import math

import numpy as np

from . import raster as orc


def l2_normalize(x, axis, eps=1e-12):
    return x / np.sqrt(np.maximum(np.sum(x * x, axis=axis, keepdims=True), eps))


def perspective(aspect_ratio, fov_y, near_clip, far_clip):
    """camera_utils.py:22-51."""
    f = 1.0 / np.tan(fov_y * (math.pi / 360.0))                      # :38
    depth_range = far_clip - near_clip                               # :39
    p_22 = -(far_clip + near_clip) / depth_range                     # :40
    p_23 = -2.0 * (far_clip * near_clip / depth_range)               # :41
    z = np.zeros_like(p_23)
    m = np.stack([f / aspect_ratio, z, z, z, z, f, z, z, z, z, p_22, p_23, z, z, -np.ones_like(p_23), z], 0)   # :44-51
    return np.transpose(m.reshape(4, 4, -1), (2, 0, 1))


def look_at(eye, center, world_up):
    """camera_utils.py:54-113."""
    B = center.shape[0]
    forward = center - eye                                                       # :74
    forward = forward / np.linalg.norm(forward, axis=1, keepdims=True)           # :75-80
    to_side = np.cross(forward, world_up)                                        # :81
    to_side = to_side / np.linalg.norm(to_side, axis=1, keepdims=True)           # :82-88
    cam_up = np.cross(to_side, forward)                                          # :89
    w_column = np.tile(np.array([[0., 0., 0., 1.]], dtype=eye.dtype), (B, 1)).reshape(B, 4, 1)   # :91-92
    view_rotation = np.stack([to_side, cam_up, -forward, np.zeros_like(to_side)], axis=1)        # :94-97
    view_rotation = np.concatenate([view_rotation, w_column], axis=2)            # :98-99
    ident = np.tile(np.eye(3, dtype=eye.dtype)[None], (B, 1, 1))
    view_translation = np.concatenate([ident, -eye[:, :, None]], 2)              # :102
    view_translation = np.concatenate([view_translation, w_column.reshape(B, 1, 4)], 1)          # :103-105
    return np.matmul(view_rotation, view_translation)                            # :106


def euler_matrices(angles):
    """camera_utils.py:116-147 (angles in radians)."""
    s, c = np.sin(angles), np.cos(angles)
    c0, c1, c2 = c[:, 0], c[:, 1], c[:, 2]
    s0, s1, s2 = s[:, 0], s[:, 1], s[:, 2]
    z, o = np.zeros_like(s0), np.ones_like(s0)
    flat = np.stack([c2 * c1, c2 * s1 * s0 - c0 * s2, s2 * s0 + c2 * c0 * s1, z,
                     c1 * s2, c2 * c0 + s2 * s1 * s0, c0 * s2 * s1 - c2 * s0, z,
                     -s1, c1 * s0, c1 * c0, z, z, z, z, o], 0)
    return np.transpose(flat.reshape(4, 4, -1), (2, 0, 1))


def rasterize_triangles(vertices, attributes, triangles, projection_matrices, image_width, image_height,
                        background_value, use_ref=None):
    """rasterize_triangles.py:30-178."""
    dt = vertices.dtype
    B, V = vertices.shape[:2]
    vh = np.concatenate([vertices, np.ones((B, V, 1), dtype=dt)], 2)                         # :73-74
    clip = np.matmul(vh, np.transpose(projection_matrices, (0, 2, 1)))                       # :78-79
    w = clip[:, :, 3:4]
    clip_w = np.maximum(np.abs(w), 1e-6) * np.sign(w)                                        # :81-84
    ndc = clip[:, :, 0:3] / clip_w                                                           # :85-86
    if use_ref is None:
        use_ref = orc.ref_available()
    fwd = orc.forward_ref if use_ref else orc.forward_port
    bary, vids = [], []
    for b in range(B):                                                                       # :88-117 (tf.map_fn)
        bc, ids, _ = fwd(ndc[b].astype(np.float32), triangles, image_width, image_height)
        bary.append(bc.reshape(-1, 3).astype(dt))
        vids.append(triangles[ids.reshape(-1)] + b * V)
    bary = np.concatenate(bary, 0)
    vids = np.concatenate(vids, 0)
    corner_attr = attributes.reshape(B * V, -1)[vids]                                        # :119-121
    corner_w = (1.0 / clip_w.reshape(-1))[vids]                                              # :125-128
    weighted = bary * corner_w                                                               # :134-135
    factor = weighted.sum(1)                                                                 # :136-137
    corrected = weighted / np.maximum(factor, 1e-6)[:, None]                                 # :139-144
    images = (corner_attr * corrected[:, :, None]).sum(1).reshape(B, image_height, image_width, -1)   # :147-152
    alphas = np.clip((2.0 * corrected).sum(1), 0.0, 1.0).reshape(B, image_height, image_width, 1)    # :157-159
    return alphas * images + (1.0 - alphas) * np.asarray(background_value, dtype=dt)         # :161-164


def phong_shader(normals, alphas, pixel_positions, light_positions, light_intensities, diffuse_colors=None,
                 camera_position=None, specular_colors=None, shininess_coefficients=None, ambient_color=None):
    """mesh_renderer.py:23-162."""
    B, H, W = normals.shape[:3]
    L = light_positions.shape[1]
    normals = normals.reshape(B, -1, 3)
    alphas = alphas.reshape(B, -1, 1)
    diffuse_colors = diffuse_colors.reshape(B, -1, 3)
    out = np.zeros((B, H * W, 3), dtype=normals.dtype)
    if ambient_color is not None:
        out = out + ambient_color[:, None, :] * diffuse_colors                               # :92-94
    pos = pixel_positions.reshape(B, -1, 3)
    to_l = l2_normalize(light_positions[:, :, None, :] - pos[:, None, :, :], 3)              # :100-102
    ndl = np.clip(np.sum(normals[:, None] * to_l, axis=3), 0.0, 1.0)                         # :106-109
    diffuse = diffuse_colors[:, None] * ndl[..., None] * light_intensities[:, :, None, :]    # :112-115
    out = out + diffuse.sum(1)                                                               # :116-118
    if camera_position is not None:
        spec = specular_colors.reshape(B, -1, 3)
        mirror = l2_normalize(2.0 * ndl[..., None] * normals[:, None] - to_l, 3)             # :123-126
        to_cam = l2_normalize(camera_position.reshape(B, 1, 3) - pos, 2)                     # :127-128
        rdc = np.sum(to_cam[:, None] * mirror, axis=3)                                       # :129-132
        rdc = np.clip(l2_normalize(rdc, 2), 0.0, 1.0)                                        # :134-136 (over pixels!)
        rdc = np.where(ndl != 0.0, rdc, np.zeros_like(rdc))                                  # :140-143
        shin = np.asarray(shininess_coefficients, dtype=normals.dtype)
        shin = shin if shin.ndim == 3 else shin.reshape(-1, 1, 1)                            # [B,H,W] | [B,1,1] | [1,1,1]
        rdc4 = rdc.reshape(B, L, H, W)                                                       # :144-146
        specularity = np.power(rdc4, shin[:, None]).reshape(B, L, H * W, 1)                  # :148-152
        out = out + (spec[:, None] * specularity * light_intensities[:, :, None, :]).sum(1)  # :153-157
    rgb = out.reshape(B, H, W, 3)
    a = alphas.reshape(B, H, W, 1)
    rgb = np.where(np.concatenate(3 * [a > 0.5], axis=3), rgb, np.zeros_like(rgb))           # :160-161
    return np.concatenate([rgb, a], axis=3)[:, ::-1]                                         # :162 (reverse rows)


def tone_mapper(image, gamma):
    """mesh_renderer.py:165-194."""
    B = image.shape[0]
    corrected = np.power(image, gamma)
    mx = corrected.reshape(B, -1).max(axis=1)
    return np.clip(corrected / mx.reshape(B, 1, 1, 1), 0.0, 1.0)


def mesh_renderer(vertices, triangles, normals, diffuse_colors, camera_position, camera_lookat, camera_up,
                  light_positions, light_intensities, image_width, image_height, specular_colors=None,
                  shininess_coefficients=None, ambient_color=None, fov_y=40.0, near_clip=0.01, far_clip=10.0,
                  use_ref=None):
    """mesh_renderer.py:197-401 (argument checks omitted)."""
    dt = vertices.dtype
    B = vertices.shape[0]

    def per_batch(x, n):
        x = np.asarray(x, dtype=dt)
        return np.tile(x.reshape(1, -1), (B, 1)) if x.shape == (n,) else x

    camera_position, camera_lookat, camera_up = (per_batch(x, 3) for x in (camera_position, camera_lookat, camera_up))
    fov_y, near_clip, far_clip = (np.broadcast_to(np.asarray(x, dtype=dt), (B,)) for x in (fov_y, near_clip, far_clip))
    per_vertex_shin = False
    if specular_colors is not None:
        shininess_coefficients = np.asarray(shininess_coefficients, dtype=dt)
        per_vertex_shin = shininess_coefficients.ndim == 2                                   # :348-359
        parts = [normals, vertices, diffuse_colors, specular_colors]
        if per_vertex_shin:
            parts.append(shininess_coefficients[:, :, None])
        attrs = np.concatenate(parts, axis=2)
    else:
        attrs = np.concatenate([normals, vertices, diffuse_colors], axis=2)                  # :361
    cam = look_at(camera_position, camera_lookat, camera_up)                                 # :363-364
    persp = perspective(image_width / image_height, fov_y, near_clip, far_clip)              # :366-367
    clip_tf = np.matmul(persp, cam)                                                          # :369
    px = rasterize_triangles(vertices, attrs, triangles, clip_tf, image_width, image_height,
                             [-1] * attrs.shape[2], use_ref)                                  # :371-373
    pn = l2_normalize(px[..., 0:3], 3)                                                       # :376
    ppos = px[..., 3:6]
    pdiff = px[..., 6:9]
    pspec, shin = None, None
    if specular_colors is not None:
        pspec = px[..., 9:12]
        shin = px[..., 12] if per_vertex_shin else shininess_coefficients.reshape(-1, 1, 1)  # :381-385
    mask = np.any(pdiff >= 0, axis=3).astype(dt)                                             # :387
    return phong_shader(pn, mask, ppos, light_positions, light_intensities, pdiff,
                        camera_position if specular_colors is not None else None, pspec, shin, ambient_color)
