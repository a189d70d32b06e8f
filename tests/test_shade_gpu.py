"""Shading kernels (csrc/shade.cu: smplb_phong_shade / smplb_phong_shade_backward) against the reference's phong_shader
graph restated in float64 torch (oracle/ref_torch.phong_shader, mesh_renderer.py:23-162) with autograd as the gradient
oracle.  Tolerances: 2e-5 max-norm on the image, 2e-4 max-norm relative on gradients (fp32 kernels vs fp64)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope='module')
def rr():
    import os
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    subprocess.check_call(['make', '-s', '-C', os.path.join(root, 'oracle'), 'port'])
    from oracle import ref_render
    return ref_render


def _scene(B, H, W, L, seed, per_pixel_shin):
    g = torch.Generator().manual_seed(seed)
    r = lambda *s: torch.randn(*s, generator=g, dtype=torch.float64)
    d = dict(normals=r(B, H, W, 3), positions=r(B, H, W, 3) * 0.5, diffuse=torch.rand(B, H, W, 3, generator=g, dtype=torch.float64),
             specular=torch.rand(B, H, W, 3, generator=g, dtype=torch.float64),
             alphas=(torch.rand(B, H, W, generator=g, dtype=torch.float64) > 0.2).double(),
             light_pos=r(B, L, 3) * 3.0 + torch.tensor([0.0, 0.0, -4.0], dtype=torch.float64),
             light_int=torch.rand(B, L, 3, generator=g, dtype=torch.float64),
             camera=r(B, 3) * 0.2 + torch.tensor([0.0, 0.0, -5.0], dtype=torch.float64),
             ambient=torch.rand(B, 3, generator=g, dtype=torch.float64) * 0.2)
    d['shin'] = (torch.rand(B, H, W, generator=g, dtype=torch.float64) * 8.0 + 1.5) if per_pixel_shin else \
        torch.tensor([[[4.0]]], dtype=torch.float64).expand(B, 1, 1).clone()
    return d


def _maxrel(a, b):
    return float((a - b).abs().max() / b.abs().max().clamp(min=1e-30))


@pytest.mark.parametrize('spec,per_pixel_shin,L', [(False, False, 2), (True, False, 1), (True, True, 3)])
def test_phong_kernels_match_reference_graph(spec, per_pixel_shin, L):
    from oracle import ref_torch as rt
    from tf_mesh_renderer import mesh_renderer as mr
    B, H, W = 2, 13, 17
    d = _scene(B, H, W, L, 7, per_pixel_shin)
    leaves64 = {k: d[k].clone().requires_grad_(True) for k in ('normals', 'positions', 'diffuse', 'specular')}
    if per_pixel_shin:
        leaves64['shin'] = d['shin'].clone().requires_grad_(True)
    n64 = rt._l2n(leaves64['normals'], 3)
    want = rt.phong_shader(n64, d['alphas'], leaves64['positions'], d['light_pos'], d['light_int'], leaves64['diffuse'],
                           d['camera'] if spec else None, leaves64['specular'] if spec else None,
                           leaves64.get('shin', d['shin']) if spec else None, d['ambient'])
    wgt = torch.randn(want.shape, generator=torch.Generator().manual_seed(3), dtype=torch.float64)
    (want * wgt).sum().backward()

    c = lambda x: x.detach().float().cuda()
    leaves = {k: c(d[k]).requires_grad_(True) for k in leaves64}
    n32 = leaves['normals'] / torch.sqrt(torch.clamp((leaves['normals'] ** 2).sum(3, keepdim=True), min=1e-12))
    got = mr.phong_shader(n32, c(d['alphas']), leaves['positions'], c(d['light_pos']), c(d['light_int']), leaves['diffuse'],
                          c(d['camera']) if spec else None, leaves['specular'] if spec else None,
                          leaves.get('shin', c(d['shin'])) if spec else None, c(d['ambient']))
    assert float((got.detach().cpu().double() - want.detach()).abs().max()) <= 2e-5
    (got * c(wgt)).sum().backward()
    names = ['normals', 'positions', 'diffuse'] + (['specular'] if spec else []) + (['shin'] if spec and per_pixel_shin else [])
    for k in names:
        assert _maxrel(leaves[k].grad.cpu().double(), leaves64[k].grad) <= 2e-4, k


def test_mesh_renderer_specular_grad_vs_reference_graph(rr):
    """The whole differentiable renderer (rasteriser + gradient kernel, interpolation, shading kernels) against the same
    pipeline with the float64 phong_shader graph after the attribute image: image and d/d vertices."""
    import render_scenes as sc
    from oracle import ref_torch as rt
    from tf_mesh_renderer import mesh_renderer as mr
    from tf_mesh_renderer import rasterize_triangles as rtri
    scene = sc.colored_cube(rr.euler_matrices, True)
    scene['image_width'], scene['image_height'] = 96, 72
    t = {k: (torch.from_numpy(v).cuda() if isinstance(v, np.ndarray) and v.dtype == np.float32 and v.ndim >= 1 else v)
         for k, v in scene.items()}
    t['vertices'] = t['vertices'].clone().requires_grad_(True)
    img = mr.mesh_renderer(**t)
    w = torch.randn(img.shape, generator=torch.Generator().manual_seed(1)).cuda()
    (img[..., :3] * w[..., :3]).sum().backward()
    g_kernel = t['vertices'].grad.clone()
    assert bool(torch.isfinite(g_kernel).all()) and float(g_kernel.abs().max()) > 0

    # the same attribute image, shaded by the float64 reference graph with autograd
    captured = {}
    orig = rtri.rasterize_triangles

    def spy(*a, **k):
        captured['px'] = orig(*a, **k)
        return captured['px']
    mr.rt.rasterize_triangles = spy
    try:
        t['vertices'].grad = None
        mr.mesh_renderer(**t)
    finally:
        mr.rt.rasterize_triangles = orig
    px = captured['px'].double()
    pn = rt._l2n(px[..., 0:3], 3)
    mask = (px[..., 6:9] >= 0).any(dim=3).double()
    shin = t['shininess_coefficients']
    shin = px[..., 12] if torch.as_tensor(shin).dim() == 2 else torch.as_tensor(shin, dtype=torch.float64).reshape(-1, 1, 1).cuda()
    want = rt.phong_shader(pn, mask, px[..., 3:6], t['light_positions'].double(), t['light_intensities'].double(),
                           px[..., 6:9], t['camera_position'].double().reshape(-1, 3).expand(px.shape[0], 3), px[..., 9:12], shin,
                           t['ambient_color'].double() if t.get('ambient_color') is not None else None)
    assert float((img.detach().double() - want.detach()).abs().max()) <= 5e-5
    (want[..., :3] * w[..., :3].double()).sum().backward()
    assert _maxrel(g_kernel.double(), t['vertices'].grad.double()) <= 5e-4
