"""Generates tests/golden/*.npz with the float64 oracle (oracle/ref_numpy.py).

These fixtures are outputs of the line-by-line restatement (the reference ships no golden vectors for this path);
the fixtures produced by the reference's own files are tests/golden/reference_*.npz (make_reference_golden.py).  Inputs: (a) the synthetic distribution of SURVEY.md §8d, (b) 16 real
SURREAL frames (pose/shape only) read from the reference's data fixture
`data/images/01_11_c0001_info.mat` when /root/reference is present (values are data, not code).

Run from the repo root:  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, '3d-pose-estimation_b200'))

from oracle import ref_numpy as rn  # noqa: E402
from smplb200 import synthetic as syn  # noqa: E402

VSUB = np.arange(0, 6890, 13)  # vertex subset stored (530 vertices) to keep the fixture small


def run(m64, beta, theta, cam_pos, cam_rot, focal):
    o = rn.smpl_forward(m64, beta, theta, tile_W=True)
    uv, px = rn.project_joints_px(o['J_transformed'], cam_pos, cam_rot, focal)
    return dict(beta=beta, theta=theta, cam_pos=cam_pos, cam_rot=cam_rot, focal=focal,
                verts_sub=o['verts'][:, VSUB], verts_sum=o['verts'].sum(axis=1),
                verts_abs_sum=np.abs(o['verts']).sum(axis=1), joints=o['joints'],
                J_transformed=o['J_transformed'], Rs=o['Rs'], A=o['A'][:, :, :3, :], uv=uv, px=px,
                v_posed_sub=o['v_posed'][:, VSUB], vsub=VSUB)


def main():
    dd = syn.make_synthetic_smpl(0)
    m64 = rn.load_model(dd, dtype=np.float64)
    inp = syn.make_inputs(8, seed=1000, dtype=np.float32)
    f64 = {k: v.astype(np.float64) for k, v in inp.items()}
    g = run(m64, f64['beta'], f64['theta'], f64['cam_pos'], f64['cam_rot'], f64['focal'])
    np.savez_compressed(os.path.join(ROOT, 'tests/golden/smpl_synth_seed0_B8.npz'), **g)

    mat = '/root/reference/data/images/01_11_c0001_info.mat'
    if os.path.exists(mat):
        import scipy.io
        d = scipy.io.loadmat(mat)
        idx = np.arange(0, 96, 6)
        theta = d['pose'].T[idx].astype(np.float64)
        beta = d['shape'].T[idx].astype(np.float64)
        cam = syn.make_inputs(len(idx), seed=7, dtype=np.float32)
        g = run(m64, beta, theta, cam['cam_pos'].astype(np.float64), cam['cam_rot'].astype(np.float64),
                cam['focal'].astype(np.float64))
        np.savez_compressed(os.path.join(ROOT, 'tests/golden/smpl_surreal16.npz'), **g)

    # known-answer vector derived from the only example in the reference (project.py:92-97)
    pts = np.array([[0, 0, 0], [1, 1, 5]], dtype=np.float64)
    cp = np.array([[0, 0, -2], [0, 0, -2]], dtype=np.float64)
    ca = np.array([[0, np.pi / 6, 0], [0, np.pi / 6, 0]], dtype=np.float64)
    fl = np.array([0.1, 0.1])
    np.savez(os.path.join(ROOT, 'tests/golden/project_kat.npz'), pts=pts, cam_pos=cp, cam_angles=ca,
             focal=fl, uv=rn.project(pts, cp, ca, fl))


if __name__ == '__main__':
    main()
