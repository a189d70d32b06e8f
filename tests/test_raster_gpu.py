"""GPU parity of the rasteriser kernels (csrc/raster.cu) through the C ABI against the pinned oracle
(oracle/raster_port.c == the reference's C++, tests/test_raster_oracle.py): ids, barycentrics and depth
bit-exact; gradient within 1e-4 (float summation order is the only difference)."""
import os
import subprocess

import numpy as np
import pytest

torch = pytest.importorskip('torch')
pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G = np.load(os.path.join(ROOT, 'tests', 'golden', 'raster_golden.npz'))


@pytest.fixture(scope='module')
def orc():
    subprocess.check_call(['make', '-s', '-C', os.path.join(ROOT, 'oracle'), 'port'])
    from oracle import raster
    return raster


def gpu_forward(ndc, tri, W, H):
    from smplb200.raster import Mesh
    ndc = np.asarray(ndc, np.float32)
    if ndc.ndim == 2:
        ndc = ndc[None]
    m = Mesh(tri, ndc.shape[1], 0)
    bary, ids, z = m.rasterize(torch.from_numpy(ndc).cuda(), W, H)
    torch.cuda.synchronize()
    return m, bary, ids, z


def assert_bit_exact(got, want):
    for g, w in zip(got, want):
        g = g.cpu().numpy()
        assert g.shape == w.shape
        assert np.array_equal(g.view(np.uint32), w.view(np.uint32)), (np.argwhere(g != w)[:5])


@pytest.mark.parametrize('name', ['tri', 'tet'])
def test_reference_golden_scenes_bit_exact(orc, name):
    v, t = G[name + '_vertices'], G[name + '_triangles']
    _, bary, ids, z = gpu_forward(v, t, 640, 480)
    assert_bit_exact((bary[0], ids[0], z[0]), orc.forward_port(v, t, 640, 480))
    img = (bary[0].cpu().numpy() * np.float32(255)).astype(np.uint8)
    assert np.abs(img.astype(int) - G[name + '_png'][..., :3].astype(int)).max() <= 1   # the reference's PNG


def test_committed_reference_outputs_bit_exact():
    W, H = int(G['rnd_W']), int(G['rnd_H'])
    _, bary, ids, z = gpu_forward(G['rnd_ndc'], G['rnd_triangles'], W, H)
    assert_bit_exact((bary[0], ids[0], z[0]), (G['rnd_bary'], G['rnd_ids'], G['rnd_z']))


@pytest.mark.parametrize('seed,V,T,W,H,B', [(0, 300, 700, 97, 61, 3), (1, 6890, 13776, 320, 240, 2), (2, 50, 40, 640, 480, 1)])
def test_random_meshes_bit_exact_batched(orc, seed, V, T, W, H, B):
    rng = np.random.default_rng(seed)
    spread = 0.05 if V > 1000 else 1.0       # SMPL-sized case: small triangles; others: huge overlapping ones
    ndc = rng.uniform(-1.2, 1.2, size=(B, V, 3)).astype(np.float32)
    tri = rng.integers(0, V, size=(T, 3)).astype(np.int32)
    if V > 1000:
        c = rng.uniform(-1.0, 1.0, size=(B, T, 1, 3))
        pts = (c + rng.normal(0, spread, size=(B, T, 3, 3))).astype(np.float32)
        V = 3 * T
        ndc = pts.reshape(B, V, 3)
        tri = np.arange(V, dtype=np.int32).reshape(T, 3)
    _, bary, ids, z = gpu_forward(ndc, tri, W, H)
    for b in range(B):
        assert_bit_exact((bary[b], ids[b], z[b]), orc.forward_port(ndc[b], tri, W, H))


def test_equal_depth_ties_go_to_the_later_triangle(orc):
    v = np.array([[-1, -1, 0.25], [1, -1, 0.25], [1, 1, 0.25], [-1, 1, 0.25]], np.float32)
    tri = np.array([[0, 1, 2], [0, 2, 3], [0, 1, 2], [2, 1, 0]], np.int32)
    _, bary, ids, z = gpu_forward(v, tri, 64, 48)
    want = orc.forward_port(v, tri, 64, 48)
    assert_bit_exact((bary[0], ids[0], z[0]), want)
    assert set(np.unique(want[1])) <= {1, 3}


def test_edge_cases(orc):
    # far-away / clipped / behind-the-plane vertices, fixed-point overflow, empty triangle list
    v = np.array([[-0.5, -0.5, 1.5], [0.0, 0.5, 1.5], [0.5, -0.5, 1.5], [300.0, 2.0, 0.1], [-250.0, -3.0, 0.2],
                  [0.1, 4e4, 0.0], [0.0, 0.0, -1.0], [0.3, 0.2, 1.0], [-0.4, 0.6, -1.0]], np.float32)
    tri = np.array([[0, 1, 2], [3, 4, 5], [6, 7, 8], [3, 6, 8], [5, 7, 4]], np.int32)
    _, bary, ids, z = gpu_forward(v, tri, 80, 60)
    assert_bit_exact((bary[0], ids[0], z[0]), orc.forward_port(v, tri, 80, 60))
    _, bary, ids, z = gpu_forward(v, np.zeros((0, 3), np.int32), 16, 8)
    assert not bary.any() and not ids.any() and bool((z == 1).all())


def test_pixel_on_edge_and_image_border(orc):
    v = np.array([0.0, -1.0, 0.5, 0.0, 1.0, 0.5, 0.5, -1.0, 0.5], np.float32).reshape(3, 3)
    for tri in ([[0, 1, 2]], [[2, 1, 0]]):
        _, bary, _, _ = gpu_forward(v, np.array(tri, np.int32), 641, 480)
        assert abs(float(bary[0, 5, 320].sum()) - 1.0) < 1e-6
    v = np.array([-1, -1, 0, 1, -1, 0, 1, 1, 0, -1, 1, 0], np.float32).reshape(4, 3)
    _, bary, _, _ = gpu_forward(v, np.array([[0, 1, 2], [0, 2, 3]], np.int32), 640, 480)
    for x, y in ((0, 0), (639, 0), (639, 479), (0, 479)):
        assert abs(float(bary[0, y, x].sum()) - 1.0) < 1e-6


@pytest.mark.parametrize('seed', [0, 1])
def test_gradient_vs_oracle(orc, seed):
    rng = np.random.default_rng(seed)
    B, T, W, H = 2, 1500, 160, 120
    c = rng.uniform(-0.9, 0.9, size=(B, T, 1, 3))
    ndc = (c + rng.normal(0, 0.06, size=(B, T, 3, 3))).astype(np.float32).reshape(B, 3 * T, 3)
    # share vertices between triangles so the per-vertex gather sums several corners
    tri = np.arange(3 * T, dtype=np.int32).reshape(T, 3)
    tri[1::2, 0] = tri[0::2, 1]
    m, bary, ids, z = gpu_forward(ndc, tri, W, H)
    gb = rng.normal(size=(B, H, W, 3)).astype(np.float32)
    g = m.rasterize_backward(torch.from_numpy(ndc).cuda(), bary, ids, torch.from_numpy(gb).cuda(), W, H).cpu().numpy()
    for b in range(B):
        want = orc.backward_port(ndc[b], tri, bary[b].cpu().numpy(), ids[b].cpu().numpy(), gb[b], W, H)
        assert np.abs(g[b] - want).max() <= 1e-4 * np.abs(want).max()
        assert np.abs(want).max() > 0


def test_drop_in_rasterize_triangles_autograd(orc):
    """tf_mesh_renderer.rasterize_triangles drop-in: identity projection, attributes = barycentric one-hots ->
    the image equals the oracle's barycentrics where covered; gradient reaches vertices and attributes."""
    from tf_mesh_renderer.rasterize_triangles import rasterize_triangles
    v = torch.tensor(G['tri_vertices'][None], device='cuda', requires_grad=True)
    attr = torch.eye(3, device='cuda')[None].clone().requires_grad_(True)
    proj = torch.eye(4, device='cuda')[None]
    img = rasterize_triangles(v, attr, G['tri_triangles'], proj, 64, 48, torch.zeros(3))
    bary, _, _ = orc.forward_port(G['tri_vertices'], G['tri_triangles'], 64, 48)
    assert img.shape == (1, 48, 64, 3)
    assert np.abs(img[0].detach().cpu().numpy() - bary).max() < 1e-5
    (img * torch.randn_like(img)).sum().backward()
    assert v.grad is not None and float(v.grad.abs().max()) > 0 and float(attr.grad.abs().max()) > 0


@pytest.mark.parametrize('B,lon,ring', [(3, 12, 10), (5, 82, 84)])
def test_vertex_normals_vs_oracle(B, lon, ring):
    """smplb_mesh_normals vs the float64 restatement of normals_from_mesh (utils.py:92-127), SMPL-sized mesh included."""
    from oracle import ref_numpy as rn
    from smplb200 import synthetic as syn
    from smplb200.raster import Mesh
    v, tri = syn.uv_sphere_mesh(lon, ring)
    rng = np.random.default_rng(lon)
    scale = rng.uniform(0.2, 1.0, size=(B, 1, 3))
    verts = (v[None] * scale + rng.normal(0, 0.002, size=(B,) + v.shape) + rng.normal(0, 0.3, size=(B, 1, 3))).astype(np.float32)
    verts[B - 1] *= -1.0          # inside-out sample: the outward test must flip it
    m = Mesh(tri, v.shape[0], 0)
    got = m.normals(torch.from_numpy(verts).cuda()).cpu().numpy()
    want = rn.normals_from_mesh(verts.astype(np.float64), tri, rn.vertex_faces_from_face_verts(tri))
    assert np.abs(got - want).max() <= 1e-5
    assert np.allclose(np.linalg.norm(got, axis=2), 1.0, atol=1e-5)
