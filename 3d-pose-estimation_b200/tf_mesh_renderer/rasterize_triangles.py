"""Drop-in for deps/tf_mesh_renderer/rasterize_triangles.py on the sm_100a rasteriser kernels.

Same call: rasterize_triangles(vertices, attributes, triangles, projection_matrices, image_width,
image_height, background_value) -> [batch, H, W, attribute_count].  The rasterisation itself
(RasterizeTriangles / RasterizeTrianglesGrad custom ops of the reference) runs in csrc/raster.cu for the whole
batch at once; the clip-space transform and the perspective-correct attribute interpolation around it are the
reference's own element-wise graph (rasterize_triangles.py:73-178) restated on torch CUDA tensors, so autograd
reaches vertices and attributes exactly as TF's does."""
import torch

from smplb200.raster import Mesh, rasterize_ndc

_MINIMUM_REWEIGHTING_THRESHOLD = 1e-6
_MINIMUM_PERSPECTIVE_DIVIDE_THRESHOLD = 1e-6
_mesh_cache = {}


def _mesh_for(triangles, vertex_count, device):
    if isinstance(triangles, Mesh):
        return triangles
    t = torch.as_tensor(triangles).to(torch.int32).cpu().contiguous()
    key = (t.numpy().tobytes(), vertex_count, str(device))
    m = _mesh_cache.get(key)
    if m is None:
        m = _mesh_cache[key] = Mesh(t, vertex_count, device)
    return m


def rasterize_triangles(vertices, attributes, triangles, projection_matrices, image_width, image_height,
                        background_value):
    if not image_width > 0:
        raise ValueError('Image width must be > 0.')
    if not image_height > 0:
        raise ValueError('Image height must be > 0.')
    if len(vertices.shape) != 3:
        raise ValueError('The vertex buffer must be 3D.')
    dev = vertices.device if vertices.is_cuda else torch.device('cuda', torch.cuda.current_device())
    vertices = vertices.to(dev, torch.float32)
    attributes = attributes.to(dev, torch.float32)
    projection_matrices = projection_matrices.to(dev, torch.float32)
    background_value = torch.as_tensor(background_value, dtype=torch.float32, device=dev)
    B, V = vertices.shape[0], vertices.shape[1]
    mesh = _mesh_for(triangles, V, dev)

    # rasterize_triangles.py:73-86: homogeneous clip space, guarded perspective divide
    vh = torch.cat([vertices, torch.ones((B, V, 1), dtype=torch.float32, device=dev)], 2)
    clip = torch.matmul(vh, projection_matrices.transpose(1, 2))
    w = clip[:, :, 3:4]
    clip_w = torch.clamp(w.abs(), min=_MINIMUM_PERSPECTIVE_DIVIDE_THRESHOLD) * torch.sign(w)
    ndc = clip[:, :, 0:3] / clip_w

    bary, ids, _ = rasterize_ndc(ndc.contiguous(), mesh, image_width, image_height)   # :88-117 (whole batch)

    # :119-152 perspective-correct barycentrics and attribute interpolation
    vids = mesh.triangles[ids.reshape(B, -1).long()] + (torch.arange(B, device=dev) * V).view(B, 1, 1)
    vids = vids.reshape(-1, 3)
    bary = bary.reshape(-1, 3)
    corner_attr = attributes.reshape(B * V, -1)[vids]
    corner_w = torch.reciprocal(clip_w.reshape(-1))[vids]
    weighted = bary * corner_w
    factor = weighted.sum(1)
    corrected = weighted / torch.clamp(factor, min=_MINIMUM_REWEIGHTING_THRESHOLD).unsqueeze(1)
    images = (corner_attr * corrected.unsqueeze(2)).sum(1).reshape(B, image_height, image_width, -1)
    # :157-178 alpha from the barycentric sum, background compositing
    alphas = torch.clamp((2.0 * corrected).sum(1), 0.0, 1.0).reshape(B, image_height, image_width, 1)
    return alphas * images + (1.0 - alphas) * background_value
