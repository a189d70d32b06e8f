"""Drop-in for deps/tf_mesh_renderer/mesh_renderer.py on the sm_100a rasteriser kernels.

mesh_renderer(...) keeps the reference signature (mesh_renderer.py:197-213).  Two paths, same results:
 * no specular terms and no gradient wanted (the way pose_3d renders, utils.py:45-50): ONE fused pass in
   csrc/raster.cu (`smplb_render_diffuse`: clip transform, rasterisation, perspective-correct interpolation and
   diffuse Phong shading; only RGBA reaches HBM);
 * otherwise rasterize_triangles (kernel rasteriser + its gradient kernel; the attribute interpolation around it is the
   reference's element-wise graph on torch tensors) followed by the shading kernels of csrc/shade.cu: specular term,
   forward and adjoint (differentiable wrt vertices, normals and colours like the TF graph).
"""
import torch

from smplb200 import shade
from smplb200.raster import Mesh

from . import camera_utils
from . import rasterize_triangles as rt


def phong_shader(normals, alphas, pixel_positions, light_positions, light_intensities, diffuse_colors=None,
                 camera_position=None, specular_colors=None, shininess_coefficients=None, ambient_color=None):
    """mesh_renderer.py:23-162 ([B,H,W,*] buffers -> lit RGBA [B,H,W,4], rows reversed) on the shading kernels of
    csrc/shade.cu (forward and adjoint).  The separate buffers of the reference signature are packed into the
    interleaved attribute image the kernels read; mesh_renderer below hands over its attribute image as it is."""
    B, H, W = normals.shape[:3]
    dev = normals.device
    parts = [normals.reshape(B, H, W, 3), pixel_positions.reshape(B, H, W, 3), diffuse_colors.reshape(B, H, W, 3)]
    off_spec = off_shin = -1
    shin_img = None
    if camera_position is not None:
        parts.append(specular_colors.reshape(B, H, W, 3))
        off_spec = 9
        shin = torch.as_tensor(shininess_coefficients, dtype=torch.float32, device=dev)
        if shin.dim() == 3 and tuple(shin.shape[1:]) == (H, W):
            parts.append(shin.expand(B, H, W).unsqueeze(3))                      # per pixel
            off_shin = 12
        else:
            shin_img = shin.reshape(-1).expand(B).contiguous()                   # [B,1,1] | [1,1,1] | scalar
    attr = torch.cat([p.to(torch.float32) for p in parts], dim=3)
    return shade.phong_shade(attr, light_positions, light_intensities, 0, 3, 6, off_spec, off_shin,
                             alphas=alphas.reshape(B, H, W), shininess=shin_img, camera_position=camera_position,
                             ambient_color=ambient_color)


def tone_mapper(image, gamma):
    """mesh_renderer.py:165-194."""
    B = image.shape[0]
    corrected = torch.pow(image, gamma)
    mx = corrected.reshape(B, -1).max(dim=1).values
    return torch.clamp(corrected / mx.reshape(B, 1, 1, 1), 0.0, 1.0)


def mesh_renderer(vertices, triangles, normals, diffuse_colors, camera_position, camera_lookat, camera_up,
                  light_positions, light_intensities, image_width, image_height, specular_colors=None,
                  shininess_coefficients=None, ambient_color=None, fov_y=40.0, near_clip=0.01, far_clip=10.0):
    if len(vertices.shape) != 3:
        raise ValueError('Vertices must have shape [batch_size, vertex_count, 3].')
    if len(normals.shape) != 3:
        raise ValueError('Normals must have shape [batch_size, vertex_count, 3].')
    if len(light_positions.shape) != 3:
        raise ValueError('Light_positions must have shape [batch_size, light_count, 3].')
    if len(light_intensities.shape) != 3:
        raise ValueError('Light_intensities must have shape [batch_size, light_count, 3].')
    if len(diffuse_colors.shape) != 3:
        raise ValueError('vertex_diffuse_colors must have shape [batch_size, vertex_count, 3].')
    if specular_colors is not None and shininess_coefficients is None:
        raise ValueError('Specular colors were supplied without shininess coefficients.')
    if shininess_coefficients is not None and specular_colors is None:
        raise ValueError('Shininess coefficients were supplied without specular colors.')
    dev = vertices.device if vertices.is_cuda else torch.device('cuda', torch.cuda.current_device())
    B, V = vertices.shape[0], vertices.shape[1]

    def t(x):
        return torch.as_tensor(x, dtype=torch.float32).to(dev)

    def per_batch(x, n, name):
        x = t(x)
        if tuple(x.shape) == (n,):
            return x.unsqueeze(0).expand(B, n)
        if tuple(x.shape) != (B, n):
            raise ValueError('%s must have shape [batch_size, %d]' % (name, n))
        return x

    def per_batch_scalar(x, name):
        x = t(x)
        if x.dim() == 0:
            return x.expand(B)
        if tuple(x.shape) != (B,):
            raise ValueError('%s must be a float, a 0D tensor, or a 1D tensor with shape [batch_size]' % name)
        return x

    vertices, normals, diffuse_colors = t(vertices), t(normals), t(diffuse_colors)
    light_positions, light_intensities = t(light_positions), t(light_intensities)
    if ambient_color is not None:
        ambient_color = t(ambient_color)
        if tuple(ambient_color.shape) != (B, 3):
            raise ValueError('Ambient_color must have shape [batch_size, 3].')
    camera_position = per_batch(camera_position, 3, 'Camera_position')
    camera_lookat = per_batch(camera_lookat, 3, 'Camera_lookat')
    camera_up = per_batch(camera_up, 3, 'Camera_up')
    fov_y, near_clip, far_clip = (per_batch_scalar(x, n) for x, n in
                                  ((fov_y, 'Fov_y'), (near_clip, 'Near_clip'), (far_clip, 'Far_clip')))
    cam = camera_utils.look_at(camera_position, camera_lookat, camera_up)                     # :363-364
    persp = camera_utils.perspective(image_width / image_height, fov_y, near_clip, far_clip)  # :366-367
    clip_tf = torch.matmul(persp, cam)                                                        # :369
    mesh = triangles if isinstance(triangles, Mesh) else rt._mesh_for(triangles, V, dev)

    wants_grad = torch.is_grad_enabled() and any(x.requires_grad for x in (vertices, normals, diffuse_colors))
    if specular_colors is None and not wants_grad:
        return mesh.render_diffuse(vertices, normals, clip_tf, light_positions, light_intensities, image_width,
                                   image_height, diffuse=diffuse_colors, ambient=ambient_color)

    per_vertex_shin = False
    if specular_colors is not None:
        specular_colors = t(specular_colors)
        shininess_coefficients = t(shininess_coefficients)
        if len(specular_colors.shape) != 3:
            raise ValueError('The specular colors must have shape [batch_size, vertex_count, 3].')
        if shininess_coefficients.dim() > 2:
            raise ValueError('The shininess coefficients must have shape at most [batch_size, vertex_count].')
        per_vertex_shin = shininess_coefficients.dim() == 2
        parts = [normals, vertices, diffuse_colors, specular_colors]
        if per_vertex_shin:
            parts.append(shininess_coefficients.unsqueeze(2))
        attrs = torch.cat(parts, dim=2)                                                        # :348-359
    else:
        attrs = torch.cat([normals, vertices, diffuse_colors], dim=2)                          # :361
    px = rt.rasterize_triangles(vertices, attrs, mesh, clip_tf, image_width, image_height,
                                [-1.0] * attrs.shape[2])                                      # :371-373
    # :375-400: normalise the interpolated normals, alpha = any(diffuse >= 0), Phong shading -- one kernel pass over the
    # attribute image as rasterize_triangles returned it (csrc/shade.cu), differentiable through smplb_phong_shade_backward
    has_spec = specular_colors is not None
    shin_img = None
    if has_spec and not per_vertex_shin:
        shin_img = shininess_coefficients.reshape(-1).expand(B).contiguous()                   # :381-385
    return shade.phong_shade(px, light_positions, light_intensities, 0, 3, 6, 9 if has_spec else -1,
                             12 if per_vertex_shin else -1, alphas=None, normalize_normals=True, shininess=shin_img,
                             camera_position=camera_position if has_spec else None, ambient_color=ambient_color)
