"""Drop-in for deps/pose_3d/utils.py (same names, argument meaning and return shapes) on torch CUDA tensors.  Every
function goes through the C ABI of libsmplb200.so; there is no CPU path.

  render_mesh_verts_cam       utils.py:12-58    normals kernel + fused clip/raster/Phong kernel
  get_camera_normal_plane     utils.py:61-74    smplb_rodrigues
  get_2d_points_in_3d         utils.py:77-89    smplb_rodrigues
  normals_from_mesh           utils.py:92-127   smplb_mesh_normals
  vertex_faces_from_face_verts utils.py:130-149 host (NumPy, runs once per topology like the reference's)
  rotate_global_pose          utils.py:152-167  smplb_rotate_global_pose
  add_axis_angle_rotations    utils.py:170-199  smplb_add_axis_angle
The heat-map helpers (soft_argmax*, gaussian_blur, utils.py:202-270) belong to the 2D CNN stage and are out of scope.
"""
import numpy as np
import torch

import smplb200
from smplb200.raster import Mesh
from tf_mesh_renderer import mesh_renderer
from tf_mesh_renderer.rasterize_triangles import _mesh_for

from . import config


def _dev(x):
    return x.device if torch.is_tensor(x) and x.is_cuda else torch.device('cuda', torch.cuda.current_device())


def _f32(x, dev):
    return torch.as_tensor(x, dtype=torch.float32).to(dev)


def _batch3(x, B, dev):
    x = _f32(x, dev)
    return x.reshape(1, 3).expand(B, 3) if tuple(x.shape) == (3,) else x


def render_mesh_verts_cam(verts, cam_pos, cam_rot, cam_fov, faces, vert_faces=None, lights=None):
    """[B,V,3] vertices seen from (cam_pos, axis-angle cam_rot, vertical fov in degrees) -> [B,H,W,1]: mean shaded
    RGB when vert_faces and lights are given, else the silhouette (alpha).  `faces` may be a smplb200.raster.Mesh.
    vert_faces is only a switch here: the adjacency lives in the Mesh."""
    dev = _dev(verts)
    verts = _f32(verts, dev)
    B = verts.shape[0]
    mesh = _mesh_for(faces, verts.shape[1], dev)
    cam_pos, cam_rot = _batch3(cam_pos, B, dev), _batch3(cam_rot, B, dev)
    R = smplb200.rodrigues(cam_rot.contiguous(), 1)                                        # :20
    cam_lookat = R[:, :, 2] + cam_pos                                                      # :22-28  R @ (0,0,1)
    cam_up = R[:, :, 1]                                                                    # :30-34  R @ (0,1,0)
    diffuse = torch.ones_like(verts)
    shaded = vert_faces is not None and lights is not None
    if not shaded:
        lights = torch.zeros((B, 1, 3), dtype=torch.float32, device=dev)                   # :37-39
        normals = torch.zeros_like(verts)
    else:
        lights = _f32(lights, dev)
        if tuple(lights.shape) == (3,):
            lights = lights.reshape(1, 1, 3).expand(B, 1, 3)                               # :41-43
        normals = mesh.normals(verts)                                                      # :44
    img_height, img_width = config.input_img_size
    rendered = mesh_renderer.mesh_renderer(verts, mesh, normals, diffuse, cam_pos, cam_lookat, cam_up,
                                           lights.contiguous(), torch.ones_like(lights), img_width, img_height,
                                           fov_y=cam_fov, near_clip=config.fl, far_clip=100.0)   # :48-52
    if not shaded:
        return rendered[:, :, :, 3:4]                                                      # :55
    return rendered[:, :, :, :3].mean(dim=3, keepdim=True)                                 # :57


def get_camera_normal_plane(cam_pos, cam_rot):
    dev = _dev(cam_pos)
    cam_pos, cam_rot = _f32(cam_pos, dev), _f32(cam_rot, dev)
    n = smplb200.rodrigues(cam_rot.contiguous(), 1)[:, :, 2]
    n = n * torch.rsqrt(torch.clamp((n * n).sum(1, keepdim=True), min=1e-12))
    return n, (cam_pos * n).sum(1)


def get_2d_points_in_3d(points_2d, cam_rot, cam_f, img_dim):
    dev = _dev(points_2d)
    p2, cam_rot, cam_f, img_dim = (_f32(x, dev) for x in (points_2d, cam_rot, cam_f, img_dim))
    rescaled = config.ss * (p2 - img_dim / 2) / img_dim[1]
    fl = cam_f.reshape(-1, 1, 1).expand(-1, p2.shape[1], 1)
    p = torch.cat([rescaled, fl], dim=2)
    R = smplb200.rodrigues(cam_rot.contiguous(), 1)
    return torch.einsum('bij,bpj->bpi', R, p)


def normals_from_mesh(vertices, faces, vertex_faces=None):
    """faces: [T,3] array or a smplb200.raster.Mesh (build it once and reuse it).  vertex_faces is accepted for
    signature compatibility; the kernel derives the same padded adjacency from the faces."""
    dev = _dev(vertices)
    vertices = _f32(vertices, dev)
    return _mesh_for(faces, vertices.shape[1], dev).normals(vertices)


def vertex_faces_from_face_verts(faces):
    faces = np.asarray(faces)
    n_vertices = int(faces.max()) + 1
    order = np.bincount(faces.reshape(-1), minlength=n_vertices)
    out = np.full((n_vertices, int(order.max())), n_vertices, dtype=np.int64)
    fill = np.zeros(n_vertices, dtype=np.int64)
    for f, (a, b, c) in enumerate(faces):
        for v in (a, b, c):
            out[v, fill[v]] = f
            fill[v] += 1
    return out


def rotate_global_pose(thetas, zrot):
    return smplb200.rotate_global_pose(thetas, zrot)


def add_axis_angle_rotations(rv1, rv2):
    return smplb200.add_axis_angle_rotations(rv1, rv2)
