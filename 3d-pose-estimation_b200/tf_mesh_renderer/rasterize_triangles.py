"""Drop-in for deps/tf_mesh_renderer/rasterize_triangles.py on the sm_100a rasteriser kernels.

Same call: rasterize_triangles(vertices, attributes, triangles, projection_matrices, image_width,
image_height, background_value) -> [batch, H, W, attribute_count].  The rasterisation itself
(RasterizeTriangles / RasterizeTrianglesGrad custom ops of the reference) runs in csrc/raster.cu for the whole
batch at once; the clip-space transform and the perspective-correct attribute interpolation around it
(rasterize_triangles.py:73-178) are kernels with hand-written adjoints too (csrc/interp.cu), so autograd reaches
vertices and attributes exactly as TF's does and every pixel and every gradient is computed by this library."""
import torch

from smplb200.raster import Mesh, clip_transform, interpolate, rasterize_ndc

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

    # rasterize_triangles.py:73-86 (clip transform, guarded perspective divide), :88-117 (rasteriser, whole batch),
    # :119-164 (perspective-correct interpolation, alpha, background): three kernels with their adjoints
    # (csrc/interp.cu, csrc/raster.cu), chained by autograd exactly where TF's graph chains the ops
    ndc, clip_w = clip_transform(vertices.contiguous(), projection_matrices)
    bary, ids, _ = rasterize_ndc(ndc, mesh, image_width, image_height)
    if background_value.dim() == 0:
        background_value = background_value.expand(attributes.shape[2])
    return interpolate(attributes.contiguous(), clip_w, bary, ids, background_value.reshape(-1), mesh)
