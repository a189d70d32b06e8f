"""Pins the rendering oracle (oracle/ref_render.py: camera matrices, perspective-correct interpolation, Phong
shading, tone mapping around the rasteriser) against the reference's golden renders with the reference's own
acceptance rule (mesh_renderer_test.py:45-166, test_utils.py:80-121)."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
subprocess.check_call(['make', '-s', '-C', os.path.join(ROOT, 'oracle'), 'port'])
from oracle import ref_render as rr  # noqa: E402
import render_scenes as sc  # noqa: E402

G = np.load(os.path.join(ROOT, 'tests', 'golden', 'render_golden.npz'))


def test_gray_cube_matches_reference_golden():
    img = rr.mesh_renderer(**sc.gray_cube(rr.euler_matrices), use_ref=False)
    assert img.shape == (2, 480, 640, 4)
    for i in range(2):
        ok, frac = sc.images_near(img[i], G['Gray_Cube_%d' % i])
        assert ok, frac


@pytest.mark.parametrize('scalar', [False, True])
def test_colored_cube_matches_reference_golden(scalar):
    r = rr.mesh_renderer(**sc.colored_cube(rr.euler_matrices, scalar), use_ref=False)
    img = np.concatenate([rr.tone_mapper(r[..., 0:3], 0.7), r[..., 3:4]], axis=3)
    for i in range(2):
        ok, frac = sc.images_near(img[i], G['Colored_Cube_%d' % i])
        assert ok, frac


def test_camera_matrices():
    eye = np.array([[0.0, 0.0, 6.0]])
    cam = rr.look_at(eye, np.zeros((1, 3)), np.array([[0.0, 1.0, 0.0]]))
    assert np.allclose(cam[0], [[1, 0, 0, 0], [0, 1, 0, 0], [0, 0, 1, -6], [0, 0, 0, 1]])
    p = rr.perspective(4 / 3, np.array([40.0]), np.array([0.01]), np.array([10.0]))[0]
    f = 1 / np.tan(np.pi * 40 / 360)
    assert np.allclose(np.diag(p)[:2], [f * 3 / 4, f]) and p[3, 2] == -1 and p[3, 3] == 0
