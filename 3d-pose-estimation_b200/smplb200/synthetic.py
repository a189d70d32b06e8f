"""Synthetic SMPL-shaped body model + synthetic inputs.

The licensed SMPL pickle the reference expects
(`applications/train_3d_pose.py:26-28`, schema written by
`deps/tf_smpl/preprocess.py:16-24`, consumed by `deps/tf_smpl/batch_smpl.py:22-66`)
is not redistributable, so parity and benchmarks use a random model with the
same shapes, sparsity and dtypes, written in the SAME pickle schema so the
drop-in loader is exercised end to end.
"""
import pickle

import numpy as np
import scipy.sparse as sp

NV = 6890
NJ = 24
NB = 10
NP = 207
NK = 19

# kintree_table[0] of the SMPL model (`batch_smpl.py:54`); entry 0 is 2**32-1 on disk.
SMPL_PARENTS = np.array(
    [-1, 0, 0, 0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 9, 9, 12, 13, 14, 16, 17, 18, 19, 20, 21],
    dtype=np.int64)


def _sparse_regressor(rng, rows, nnz_per_row):
    """[rows, NV] scipy-sparse with `nnz_per_row` Dirichlet weights per row (rows sum to 1)."""
    data, ri, ci = [], [], []
    for r in range(rows):
        cols = rng.choice(NV, size=nnz_per_row, replace=False)
        w = rng.dirichlet(np.ones(nnz_per_row))
        data.append(w)
        ri.append(np.full(nnz_per_row, r))
        ci.append(cols)
    m = sp.csc_matrix(
        (np.concatenate(data), (np.concatenate(ri), np.concatenate(ci))), shape=(rows, NV))
    return m


def make_synthetic_smpl(seed=0, dense_weights=False):
    """Return a dict in the reference pkl schema (fp64 generated, fp32-representable values)."""
    rng = np.random.default_rng(seed)
    v_template = rng.normal(0.0, 0.3, size=(NV, 3))
    v_template[:, 1] *= 2.0
    shapedirs = rng.normal(0.0, 0.01, size=(NV, 3, NB))
    posedirs = rng.normal(0.0, 0.005, size=(NV, 3, NP))
    weights = np.zeros((NV, NJ))
    if dense_weights:
        weights = rng.dirichlet(np.ones(NJ), size=NV)
    else:
        for v in range(NV):
            js = rng.choice(NJ, size=4, replace=False)
            weights[v, js] = rng.dirichlet(np.ones(4))
    J_regressor = _sparse_regressor(rng, NJ, 32)
    cocoplus = _sparse_regressor(rng, NK, 32)
    kintree = np.zeros((2, NJ), dtype=np.uint32)
    kintree[0, 1:] = SMPL_PARENTS[1:].astype(np.uint32)
    kintree[0, 0] = np.uint32(2 ** 32 - 1)
    kintree[1] = np.arange(NJ, dtype=np.uint32)

    def f32(x):  # values exactly representable in fp32, held as fp64 like the real pkl
        return np.asarray(x, dtype=np.float32).astype(np.float64)

    # fp32-rounded LBS weights no longer sum to exactly 1; that is also true of the real model.
    J_regressor = J_regressor.astype(np.float32).astype(np.float64)
    cocoplus = cocoplus.astype(np.float32).astype(np.float64)
    return {
        'v_template': f32(v_template),
        'shapedirs': f32(shapedirs),
        'posedirs': f32(posedirs),
        'J_regressor': J_regressor,
        'weights': f32(weights),
        'kintree_table': kintree,
        'cocoplus_regressor': cocoplus,
    }


def write_pkl(path, dd):
    with open(path, 'wb') as f:
        pickle.dump(dd, f, protocol=2)


def make_inputs(B, seed=1000, dtype=np.float32):
    """Synthetic per-sample inputs: beta, theta, camera."""
    rng = np.random.default_rng(seed)
    beta = np.clip(rng.normal(0.0, 1.0, size=(B, NB)), -3.0, 3.0)
    theta = rng.normal(0.0, 0.3, size=(B, NJ * 3))
    axis = rng.normal(size=(B, 3))
    axis /= np.linalg.norm(axis, axis=1, keepdims=True)
    theta[:, :3] = axis * rng.uniform(0.0, np.pi, size=(B, 1))
    cam_pos = np.stack([rng.normal(0.0, 0.1, B), rng.normal(-0.35, 0.1, B),
                        rng.uniform(-6.0, -3.0, B)], axis=1)
    cam_rot = rng.normal(0.0, 0.05, size=(B, 3))
    focal = np.full((B,), 0.05)
    return {
        'beta': beta.astype(dtype), 'theta': theta.astype(dtype),
        'cam_pos': cam_pos.astype(dtype), 'cam_rot': cam_rot.astype(dtype),
        'focal': focal.astype(dtype),
    }


def uv_sphere_mesh(n_lon=82, n_ring=84):
    """Closed UV-sphere with exactly SMPL's counts: 82*84 + 2 = 6890 vertices, 82*83*2 + 2*82 = 13776 triangles."""
    th = np.pi * (np.arange(n_ring) + 1) / (n_ring + 1)
    ph = 2 * np.pi * np.arange(n_lon) / n_lon
    ring = np.stack([np.outer(np.sin(th), np.cos(ph)), np.outer(np.cos(th), np.ones(n_lon)),
                     np.outer(np.sin(th), np.sin(ph))], -1).reshape(-1, 3)
    v = np.concatenate([ring, [[0, 1, 0]], [[0, -1, 0]]]).astype(np.float32)
    idx = lambda r, c: r * n_lon + (c % n_lon)
    tri = []
    for r in range(n_ring - 1):
        for c in range(n_lon):
            tri += [[idx(r, c), idx(r + 1, c), idx(r, c + 1)], [idx(r, c + 1), idx(r + 1, c), idx(r + 1, c + 1)]]
    top, bot = n_lon * n_ring, n_lon * n_ring + 1
    for c in range(n_lon):
        tri += [[top, idx(0, c), idx(0, c + 1)], [bot, idx(n_ring - 1, c + 1), idx(n_ring - 1, c)]]
    return v, np.asarray(tri, np.int32)
