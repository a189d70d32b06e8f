"""SURREAL annotation loader (smplb200/surreal.py) against the arrays data_helpers.py:58-88 makes of
the reference's own sample file (tests/golden/surreal_info_c0001.npz, make_surreal_info_golden.py)."""
import os

import numpy as np
import pytest

from conftest import ROOT

G = np.load(os.path.join(ROOT, 'tests', 'golden', 'surreal_info_c0001.npz'))


def write_mat(tmp_path):
    import scipy.io
    p = str(tmp_path / 'c0001_info.mat')
    scipy.io.savemat(p, {'pose': G['raw_pose'], 'shape': G['raw_shape'], 'joints2D': G['raw_joints2D'], 'zrot': G['raw_zrot']})
    return p


def test_read_info_matches_reference_processing(tmp_path):
    from smplb200 import surreal
    ann = surreal.read_info_surreal(write_mat(tmp_path))
    for k in ('poses', 'shapes', 'joints2d', 'zrot'):
        assert ann[k].dtype == np.float32
        assert np.array_equal(ann[k], G[k].astype(np.float32)), k
    assert ann['poses'].shape == (50, 72) and ann['joints2d'].shape == (50, 24, 2)
    mask = np.zeros(100, bool)
    mask[10:30] = True
    ann = surreal.read_info_surreal(write_mat(tmp_path), mask=mask, skip=2)
    assert np.array_equal(ann['zrot'], np.squeeze(G['raw_zrot'])[10:30:2])


def test_reader_matches_reference_function_run(tmp_path):
    """tests/golden/reference_surreal.npz is what the reference's read_maps_poses_images_surreal itself returned
    (data_helpers.py:43-89, run by make_reference_golden.py) for its own annotation sample plus a synthetic heat-map
    file whose mask loads as uint8 -- so its `x[mask]` is integer indexing (frames 0/1), which the loader reproduces."""
    import scipy.io
    from smplb200 import surreal
    R = np.load(os.path.join(ROOT, 'tests', 'golden', 'reference_surreal.npz'))
    T, W = R['mask'].shape[0], int(R['img_size_x'])
    maps = str(tmp_path / 'maps.mat')
    scipy.io.savemat(maps, dict(heat_mat=np.zeros((2, W, 19, T), np.float32), mask=R['mask'], diffs=R['diffs']))
    mask, w = surreal.read_maps_mask_surreal(maps)
    assert w == W and mask.dtype == np.uint8 and mask.shape == (T,)
    with np.errstate(invalid='ignore'):
        assert np.array_equal(mask.astype(bool), (np.squeeze(R['mask']) != 0) & (np.squeeze(R['diffs']) < 250))
    ann = surreal.read_info_surreal(write_mat(tmp_path), img_size_x=w, mask=mask)
    for k in ('poses', 'shapes', 'joints2d', 'zrot'):
        assert ann[k].shape == R[k].shape, k
        assert np.array_equal(ann[k], R[k].astype(np.float32)), k
    # the selection the author intended needs a bool mask
    sel = surreal.read_info_surreal(write_mat(tmp_path), img_size_x=w, mask=mask.astype(bool))
    assert sel['poses'].shape[0] == (int(mask.astype(bool).sum()) + 1) // 2


@pytest.mark.gpu
def test_gt_batch_on_gpu(tmp_path, gpu_model, oracle64):
    from oracle import ref_numpy as rn
    from smplb200 import surreal
    ann = surreal.read_info_surreal(write_mat(tmp_path))
    gt = surreal.gt_batch(gpu_model, ann)
    th = rn.rotate_global_pose(ann['poses'].astype(np.float64), ann['zrot'].astype(np.float64))
    # rotations compared as matrices.  These frames compose to angles close to 2*pi (|c| ~ 6.1 rad), where
    # n = sin(c/2) n / sin(c/2) (utils.py:194) divides by ~0.07: the float32 formula itself is only good to ~1e-5
    # there, so the bound is the float32 restatement's own error (x4) rather than 1e-5
    R64 = rn.rodrigues_batch(th[:, :3])
    th32 = rn.rotate_global_pose(ann['poses'], ann['zrot'])
    floor = np.abs(rn.rodrigues_batch(th32[:, :3].astype(np.float64)) - R64).max()
    err = np.abs(rn.rodrigues_batch(gt['theta'][:, :3].double().cpu().numpy()) - R64).max()
    assert err < max(4 * floor, 1e-5), (err, floor)
    o = rn.smpl_forward(oracle64, ann['shapes'].astype(np.float64), gt['theta'].double().cpu().numpy(), tile_W=False)
    assert np.abs(gt['verts'].cpu().numpy() - o['verts']).max() / np.abs(o['verts']).max() < 1e-5
    assert np.abs(gt['joints3d'].cpu().numpy() - o['J_transformed']).max() / np.abs(o['J_transformed']).max() < 1e-5
    assert np.allclose(gt['joints2d_flipped'][..., 1].cpu().numpy(), 240.0 - ann['joints2d'][..., 1])
