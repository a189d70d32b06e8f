"""Drop-in for deps/tf_mesh_renderer/camera_utils.py (torch tensors as carriers; [B,4,4] matrices are host-side
plumbing for the renderer kernels, a few dozen flops per image)."""
import math

import torch


def perspective(aspect_ratio, fov_y, near_clip, far_clip):
    """camera_utils.py:22-51: [B] fov (degrees), near, far -> [B,4,4] eye space -> clip space."""
    f = 1.0 / torch.tan(fov_y * (math.pi / 360.0))
    depth_range = far_clip - near_clip
    p_22 = -(far_clip + near_clip) / depth_range
    p_23 = -2.0 * (far_clip * near_clip / depth_range)
    z, o = torch.zeros_like(p_23), torch.ones_like(p_23)
    m = torch.stack([f / aspect_ratio, z, z, z, z, f, z, z, z, z, p_22, p_23, z, z, -o, z], 0)
    return m.reshape(4, 4, -1).permute(2, 0, 1)


def look_at(eye, center, world_up):
    """camera_utils.py:54-113: gluLookAt-style world -> eye matrices [B,4,4]."""
    B = center.shape[0]
    cutoff = 1e-6
    forward = center - eye
    fn = torch.linalg.norm(forward, dim=1, keepdim=True)
    if not bool((fn > cutoff).all()):
        raise ValueError('Camera matrix is degenerate because eye and center are close.')
    forward = forward / fn
    to_side = torch.cross(forward, world_up, dim=1)
    sn = torch.linalg.norm(to_side, dim=1, keepdim=True)
    if not bool((sn > cutoff).all()):
        raise ValueError('Camera matrix is degenerate because up and gaze are close or because up is degenerate.')
    to_side = to_side / sn
    cam_up = torch.cross(to_side, forward, dim=1)
    w_col = torch.tensor([0., 0., 0., 1.], dtype=eye.dtype, device=eye.device).repeat(B, 1).reshape(B, 4, 1)
    rot = torch.cat([torch.stack([to_side, cam_up, -forward, torch.zeros_like(to_side)], dim=1), w_col], dim=2)
    ident = torch.eye(3, dtype=eye.dtype, device=eye.device).repeat(B, 1, 1)
    trans = torch.cat([torch.cat([ident, -eye.unsqueeze(2)], 2), w_col.reshape(B, 1, 4)], 1)
    return torch.matmul(rot, trans)


def euler_matrices(angles):
    """camera_utils.py:116-147: XYZ Tait-Bryan rotation, angles [B,3] in radians -> [B,4,4]."""
    s, c = torch.sin(angles), torch.cos(angles)
    c0, c1, c2 = c[:, 0], c[:, 1], c[:, 2]
    s0, s1, s2 = s[:, 0], s[:, 1], s[:, 2]
    z, o = torch.zeros_like(s0), torch.ones_like(s0)
    flat = torch.stack([c2 * c1, c2 * s1 * s0 - c0 * s2, s2 * s0 + c2 * c0 * s1, z,
                        c1 * s2, c2 * c0 + s2 * s1 * s0, c0 * s2 * s1 - c2 * s0, z,
                        -s1, c1 * s0, c1 * c0, z, z, z, z, o], 0)
    return flat.reshape(4, 4, -1).permute(2, 0, 1)
