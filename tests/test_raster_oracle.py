"""Pins the rasteriser oracle: the C restatement (oracle/raster_port.c) against
 (a) the reference's golden images (test_data/Simple_Triangle.png, Simple_Tetrahedron.png; decoded in
     tests/golden/raster_golden.npz, see make_raster_golden.py),
 (b) the reference's own C++ compiled into oracle/_ref (bit-exact, when that build exists),
 (c) the committed outputs of that C++ on a seeded random mesh (bit-exact, always),
 (d) the reference's edge cases (rasterize_triangles_impl_test.cc:103-141),
and the gradient restatement against central finite differences (rasterize_triangles_test.py:106-179 idea)."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
subprocess.check_call(['make', '-s', '-C', os.path.join(ROOT, 'oracle'), 'port'])
from oracle import raster as orc  # noqa: E402

G = np.load(os.path.join(ROOT, 'tests', 'golden', 'raster_golden.npz'))


def to_rgba(bary):
    """rasterize_triangles_impl_test.cc:35-47 (float RGB * 255 -> uint8, alpha 255)."""
    out = np.full(bary.shape[:2] + (4,), 255, np.uint8)
    out[..., :3] = (bary * np.float32(255)).astype(np.uint8)
    return out


@pytest.mark.parametrize('name', ['tri', 'tet'])
def test_port_matches_reference_golden_png(name):
    bary, ids, z = orc.forward_port(G[name + '_vertices'], G[name + '_triangles'], 640, 480)
    img = to_rgba(bary)
    ref = G[name + '_png']
    diff = np.abs(img.astype(int) - ref.astype(int))
    # the golden PNGs were written by the reference on another platform: allow 1 LSB on isolated pixels
    assert diff.max() <= 1, diff.max()
    assert (diff > 0).mean() < 1e-3


def test_port_matches_committed_reference_outputs_bit_exact():
    bary, ids, z = orc.forward_port(G['rnd_ndc'], G['rnd_triangles'], int(G['rnd_W']), int(G['rnd_H']))
    assert np.array_equal(ids, G['rnd_ids'])
    assert np.array_equal(bary.view(np.uint32), G['rnd_bary'].view(np.uint32))
    assert np.array_equal(z.view(np.uint32), G['rnd_z'].view(np.uint32))
    assert (G['rnd_bary'].sum(-1) > 0.5).mean() > 0.2      # the fixture actually covers pixels


@pytest.mark.skipif(not orc.ref_available(), reason='oracle/_ref not built (reference sources absent)')
@pytest.mark.parametrize('seed', [0, 1, 2])
def test_port_matches_reference_cpp_bit_exact(seed):
    rng = np.random.default_rng(seed)
    V, T, W, H = 300, 700, 97, 61
    ndc = rng.uniform(-1.2, 1.2, size=(V, 3)).astype(np.float32)
    tri = rng.integers(0, V, size=(T, 3)).astype(np.int32)
    tri[:, 1] = (tri[:, 0] + rng.integers(1, 9, size=T)) % V
    a = orc.forward_port(ndc, tri, W, H)
    b = orc.forward_ref(ndc, tri, W, H)
    for x, y in zip(a, b):
        assert np.array_equal(x.view(np.uint32), y.view(np.uint32))


def test_pixel_on_triangle_edge_is_covered_both_windings():
    W, H = 641, 480
    v = np.array([0.0, -1.0, 0.5, 0.0, 1.0, 0.5, 0.5, -1.0, 0.5], np.float32).reshape(3, 3)
    for tri in ([[0, 1, 2]], [[2, 1, 0]]):
        bary, _, _ = orc.forward_port(v, np.array(tri, np.int32), W, H)
        assert abs(bary[5, W // 2].sum() - 1.0) < 1e-6


def test_covers_edge_pixels_of_image():
    W, H = 640, 480
    v = np.array([-1, -1, 0, 1, -1, 0, 1, 1, 0, -1, 1, 0], np.float32).reshape(4, 3)
    bary, _, _ = orc.forward_port(v, np.array([[0, 1, 2], [0, 2, 3]], np.int32), W, H)
    for x, y in ((0, 0), (W - 1, 0), (W - 1, H - 1), (0, H - 1)):
        assert abs(bary[y, x].sum() - 1.0) < 1e-6


def test_empty_and_clipped():
    v = np.array([[-0.5, -0.5, 1.5], [0.0, 0.5, 1.5], [0.5, -0.5, 1.5]], np.float32)   # beyond the far plane
    bary, ids, z = orc.forward_port(v, np.array([[0, 1, 2]], np.int32), 32, 24)
    assert not bary.any() and not ids.any() and (z == 1).all()
    bary, ids, z = orc.forward_port(v, np.zeros((0, 3), np.int32), 32, 24)
    assert not bary.any() and (z == 1).all()


def test_gradient_restatement_vs_finite_differences():
    """df/dvertices for f = sum(w * bary): interior pixels of one big triangle, central differences in float64
    of the barycentric formula (rasterize_triangles_impl.cc:124-132) for the same pixel set."""
    W, H = 48, 36
    v = np.array([[-0.8, -0.7, 0.2], [0.1, 0.8, 0.1], [0.7, -0.6, 0.3]], np.float32)
    tri = np.array([[0, 1, 2]], np.int32)
    bary, ids, _ = orc.forward_port(v, tri, W, H)
    rng = np.random.default_rng(3)
    w = rng.normal(size=(H, W, 3)).astype(np.float32)
    g = orc.backward_port(v, tri, bary, ids, w, W, H)
    mask = bary.sum(-1) > 0.9

    def f(vv):
        vx = (vv[:, 0] + 1.0) * 0.5 * W
        vy = (vv[:, 1] + 1.0) * 0.5 * H
        iy, ix = np.nonzero(mask)
        px, py = ix + 0.5, iy + 0.5
        area = (vx[2] - vx[0]) * (vy[1] - vy[0]) - (vy[2] - vy[0]) * (vx[1] - vx[0])
        b0 = ((px - vx[1]) * (vy[2] - vy[1]) - (py - vy[1]) * (vx[2] - vx[1])) / area
        b1 = ((px - vx[2]) * (vy[0] - vy[2]) - (py - vy[2]) * (vx[0] - vx[2])) / area
        b2 = 1 - b0 - b1
        ww = w[iy, ix].astype(np.float64)
        return (ww[:, 0] * b0 + ww[:, 1] * b1 + ww[:, 2] * b2).sum()

    fd = np.zeros((3, 3))
    for i in range(3):
        for c in range(2):
            d = np.zeros((3, 3))
            d[i, c] = 1e-5
            fd[i, c] = (f(v.astype(np.float64) + d) - f(v.astype(np.float64) - d)) / 2e-5
    assert np.abs(g - fd).max() <= 1e-3 * np.abs(fd).max(), (g, fd)
    assert not g[:, 2].any()


def test_normals_restatement_on_sphere_and_pad_quirk():
    """normals_from_mesh restatement (oracle/ref_numpy.py): unit length, outward on a sphere, and the reference's
    pad-index behaviour (pad = n_vertices is a valid face id when faces > vertices)."""
    from oracle import ref_numpy as rn
    from smplb200 import synthetic as syn
    v, tri = syn.uv_sphere_mesh(12, 10)
    vf = rn.vertex_faces_from_face_verts(tri)
    assert vf.shape[0] == v.shape[0] and (vf == v.shape[0]).any() and tri.shape[0] > v.shape[0]
    n = rn.normals_from_mesh(v[None].astype(np.float64), tri, vf)[0]
    assert np.allclose(np.linalg.norm(n, axis=1), 1.0, atol=1e-9)
    # vertices of maximal order (the two poles) get no pad contribution and point outward on the unit sphere
    full = (vf != v.shape[0]).all(1)
    assert full.sum() == 2 and ((n[full] * v[full]).sum(1) > 0.8).all()
    # every other vertex also sums (max_order - order) copies of face n_vertices' normal (utils.py:145-147)
    fn = rn.l2_normalize(np.cross(v[tri[:, 1]] - v[tri[:, 0]], v[tri[:, 2]] - v[tri[:, 0]])[None].astype(np.float64), 1)[0]
    k = 5
    own = [t for t in range(tri.shape[0]) if k in tri[t]]
    raw = fn[own].sum(0) + (vf.shape[1] - len(own)) * fn[v.shape[0]]
    raw /= np.linalg.norm(raw)
    assert np.allclose(np.abs(raw), np.abs(n[k]), atol=1e-9)
