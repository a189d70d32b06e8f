"""H36M TFRecord reader (smplb200/h36m.py) and the heat-map / frame half of the SURREAL reader against
what the reference's own functions returned (tests/golden/reference_{h36m,surreal}.npz, made by
make_reference_golden.py running deps/pose_3d/data_helpers.py on the TensorFlow stand-in, whose Example parsing is the
official protobuf runtime -- not the hand-written wire-format parser under test)."""
import os

import numpy as np
import pytest

from conftest import ROOT

GOLD = os.path.join(ROOT, 'tests', 'golden')


def test_crc32c_known_answers():
    from smplb200 import h36m
    assert h36m.crc32c(b'123456789') == 0xE3069283          # the standard CRC-32C check value
    assert h36m.crc32c(b'') == 0
    assert h36m.crc32c(bytes(32)) == 0x8A9136AA             # RFC 3720 B.4: 32 bytes of zeros
    assert h36m.crc32c(bytes([0xFF] * 32)) == 0x62A8AB43    # RFC 3720 B.4: 32 bytes of ones


def test_tfrecord_reader_matches_reference_function(tmp_path):
    from smplb200 import h36m
    R = np.load(os.path.join(GOLD, 'reference_h36m.npz'))
    path = str(tmp_path / 'h36m.tfrecord')
    R['tfrecord'].tofile(path)
    recs = list(h36m.tfrecord_iter(path))
    assert len(recs) == R['poses'].shape[0]
    for i, rec in enumerate(recs):
        img, p, s, j = h36m.read_tfrecord_h36m(rec)
        assert p.dtype == s.dtype == j.dtype == np.float32
        assert np.array_equal(p, R['poses'][i]) and np.array_equal(s, R['shapes'][i])
        assert j.shape == (14, 2) and np.array_equal(j, R['joints2d'][i])
        assert np.array_equal(img, R['imgs'][i])            # same JPEG decoder (PIL) as the golden run
    ann = h36m.read_annotations_h36m(path)
    assert np.array_equal(ann['poses'], R['poses']) and np.array_equal(ann['joints2d'], R['joints2d'])
    # framing errors are reported, not skipped
    bad = R['tfrecord'].copy()
    bad[40] ^= 0x10
    bad.tofile(path)
    with pytest.raises(ValueError):
        list(h36m.tfrecord_iter(path))
    # a record without one of the FixedLenFeature keys is an error like tf.parse_single_example's
    with pytest.raises(ValueError):
        h36m.read_tfrecord_h36m(b'')


def test_h36m_heatmaps_and_concat(tmp_path):
    import scipy.io
    from smplb200 import h36m
    R = np.load(os.path.join(GOLD, 'reference_h36m.npz'))
    maps = str(tmp_path / 'maps.mat')
    scipy.io.savemat(maps, dict(heat_mat=R['heat_mat']))
    heat = h36m.read_maps_h36m(maps)
    assert np.array_equal(heat, R['heatmaps'])
    c, p, s, j, z = h36m.concat_img_heatmaps_h36m(heat, R['imgs'], R['poses'], R['shapes'], R['joints2d'])
    assert c.shape == heat.shape[:3] + (14 + 3,) and z == 0.0
    assert np.array_equal(c[..., 14:], R['imgs'].astype(np.float32))


def test_surreal_heatmap_frame_half_matches_reference_function(tmp_path):
    import scipy.io
    from smplb200 import surreal
    R = np.load(os.path.join(GOLD, 'reference_surreal.npz'))
    maps = str(tmp_path / 'maps.mat')
    scipy.io.savemat(maps, dict(heat_mat=R['heat_mat'], mask=R['mask'], diffs=R['diffs']))
    frames = tmp_path / 'frames'
    frames.mkdir()
    off = 0
    for t, n in enumerate(R['jpeg_len']):
        (frames / ('f%04d.jpg' % t)).write_bytes(R['jpeg_blob'][off:off + n].tobytes())
        off += int(n)
    mask, w = surreal.read_maps_mask_surreal(maps)
    concat = surreal.read_maps_frames_surreal(maps, str(frames).encode('utf-8'), mask=mask)
    assert concat.shape == R['concat'].shape and concat.dtype == np.float32
    assert np.array_equal(concat[..., :14], R['concat'][..., :14])                      # heat maps: exact
    assert np.abs(concat[..., 14:] - R['concat'][..., 14:]).max() <= 1e-6               # frames: min-max normalised


@pytest.mark.gpu
def test_h36m_batch_on_gpu(tmp_path, gpu_model, oracle64):
    from oracle import ref_numpy as rn
    from smplb200 import h36m
    R = np.load(os.path.join(GOLD, 'reference_h36m.npz'))
    path = str(tmp_path / 'h36m.tfrecord')
    R['tfrecord'].tofile(path)
    ann = h36m.read_annotations_h36m(path)
    gt = h36m.h36m_batch(gpu_model, ann)
    o = rn.smpl_forward(oracle64, ann['shapes'].astype(np.float64), ann['poses'].astype(np.float64), tile_W=False)
    assert np.abs(gt['verts'].cpu().numpy() - o['verts']).max() / np.abs(o['verts']).max() < 1e-5
    assert np.abs(gt['keypoints'].cpu().numpy() - o['joints']).max() / np.abs(o['joints']).max() < 1e-5
