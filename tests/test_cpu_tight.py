"""The tight C + OpenMP CPU version (oracle/smpl_omp.c) against the float64 oracle: it is timed by
bench.py as cpu_baseline.tight, so it has to compute the same thing."""
import numpy as np
import pytest

from oracle import ref_numpy as rn
from oracle import smpl_omp
from smplb200 import synthetic as syn


@pytest.mark.parametrize('B', [1, 7, 19])
def test_tight_cpu_matches_oracle(B):
    dd = syn.make_synthetic_smpl(0)
    m32, m64 = rn.load_model(dd, dtype=np.float32), rn.load_model(dd, dtype=np.float64)
    inp = syn.make_inputs(B, seed=5)
    tm = smpl_omp.TightModel(m32)
    got = tm.forward(inp['beta'], inp['theta'], inp['cam_pos'], inp['cam_rot'], inp['focal'], rn.CFG.ss, 320.0, 240.0)
    d = {k: v.astype(np.float64) for k, v in inp.items()}
    o = rn.smpl_forward(m64, d['beta'], d['theta'], tile_W=False)
    _, px = rn.project_joints_px(o['J_transformed'], d['cam_pos'], d['cam_rot'], d['focal'])
    for k, want in (('verts', o['verts']), ('J_transformed', o['J_transformed']), ('joints', o['joints'])):
        assert np.max(np.abs(got[k] - want)) <= 1e-5, k          # fp32 against fp64, the tolerance of the GPU tests
    assert np.max(np.abs(got['px'] - px)) <= 1e-4 * np.max(np.abs(px))
    # without the vertex output (per-thread scratch) and single-threaded: same numbers
    g2 = tm.forward(inp['beta'], inp['theta'], inp['cam_pos'], inp['cam_rot'], inp['focal'], rn.CFG.ss, 320.0, 240.0,
                    want_verts=False, nthreads=1)
    assert g2['threads'] == 1
    assert np.array_equal(g2['joints'], got['joints']) and np.array_equal(g2['px'], got['px'])
