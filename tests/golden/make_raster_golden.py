"""Generates tests/golden/raster_golden.npz (run in the build container, where /root/reference exists).

 * the reference's own golden images for the rasteriser (deps/tf_mesh_renderer/test_data/Simple_Triangle.png,
   Simple_Tetrahedron.png: barycentrics * 255 as uint8 RGBA at 640x480, written by
   kernels/rasterize_triangles_impl_test.cc:35-47,82-101) decoded to arrays, with the inputs of those tests;
 * outputs of the reference's own C++ (oracle/_ref) on a seeded SMPL-sized random mesh (bit-exact fixture)."""
import os
import sys

import numpy as np
from PIL import Image

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import raster as orc  # noqa: E402

TD = '/root/reference/deps/tf_mesh_renderer/test_data'
out = {}
out['tri_vertices'] = np.array([-0.5, -0.5, 0.8, 0.0, 0.5, 0.3, 0.5, -0.5, 0.3], np.float32).reshape(3, 3)
out['tri_triangles'] = np.array([[0, 1, 2]], np.int32)
out['tri_png'] = np.asarray(Image.open(os.path.join(TD, 'Simple_Triangle.png')).convert('RGBA'))
out['tet_vertices'] = np.array([-0.5, -0.5, 0.8, 0.0, 0.5, 0.3, 0.5, -0.5, 0.3, 0.0, 0.0, 0.0], np.float32).reshape(4, 3)
out['tet_triangles'] = np.array([0, 2, 1, 0, 1, 3, 1, 2, 3, 2, 0, 3], np.int32).reshape(4, 3)
out['tet_png'] = np.asarray(Image.open(os.path.join(TD, 'Simple_Tetrahedron.png')).convert('RGBA'))

# seeded random "mesh": 400 vertices, 900 random triangles, 160x120 (sparse storage of the covered pixels)
rng = np.random.default_rng(7)
V, T, W, H = 400, 900, 160, 120
ndc = rng.uniform(-1.1, 1.1, size=(V, 3)).astype(np.float32)
ndc[:, 2] = rng.uniform(-1.2, 1.2, size=V).astype(np.float32)
base = rng.integers(0, V, size=T)
tri = np.stack([base, (base + rng.integers(1, 12, size=T)) % V, (base + rng.integers(12, 24, size=T)) % V], 1).astype(np.int32)
bary, ids, z = orc.forward_ref(ndc, tri, W, H)
out.update(rnd_ndc=ndc, rnd_triangles=tri, rnd_W=W, rnd_H=H, rnd_bary=bary, rnd_ids=ids, rnd_z=z)
np.savez_compressed(os.path.join(ROOT, 'tests', 'golden', 'raster_golden.npz'), **out)
print({k: getattr(v, 'shape', v) for k, v in out.items()})
