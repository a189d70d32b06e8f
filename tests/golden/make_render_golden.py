"""Generates tests/golden/render_golden.npz: the reference's golden renders (deps/tf_mesh_renderer/test_data/
Gray_Cube_{0,1}.png, Colored_Cube_{0,1}.png, produced by mesh_renderer_test.py:45-166) decoded to uint8 arrays.
Run in the build container (needs /root/reference)."""
import os

import numpy as np
from PIL import Image

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
TD = '/root/reference/deps/tf_mesh_renderer/test_data'
out = {}
for n in ('Gray_Cube_0', 'Gray_Cube_1', 'Colored_Cube_0', 'Colored_Cube_1'):
    out[n] = np.asarray(Image.open(os.path.join(TD, n + '.png')))
    print(n, out[n].shape, out[n].dtype)
np.savez_compressed(os.path.join(ROOT, 'tests', 'golden', 'render_golden.npz'), **out)
