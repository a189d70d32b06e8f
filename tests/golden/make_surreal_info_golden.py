"""Generates tests/golden/surreal_info_c0001.npz from the reference's own SURREAL annotation sample
(/root/reference/data/images/01_11_c0001_info.mat, 100 frames): the raw arrays the loader consumes (so the test can
re-create a .mat with scipy.io.savemat) and what data_helpers.py:58-88 makes of them.  Build container only."""
import os

import numpy as np
import scipy.io

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
d = scipy.io.loadmat('/root/reference/data/images/01_11_c0001_info.mat')
raw = {k: d[k] for k in ('pose', 'shape', 'joints2D', 'zrot')}
poses = np.transpose(d['pose'], (1, 0))                      # data_helpers.py:62
shapes = np.transpose(d['shape'], (1, 0))                    # :63
j2d = np.transpose(d['joints2D'], (2, 1, 0)).copy()          # :65
j2d[:, :, 0] = 320 - j2d[:, :, 0]                            # :67 (heat maps are 320 wide)
zrot = np.squeeze(np.array(d['zrot']))                       # :68
skip = 2                                                     # :84-87
np.savez_compressed(os.path.join(ROOT, 'tests', 'golden', 'surreal_info_c0001.npz'),
                    raw_pose=raw['pose'], raw_shape=raw['shape'], raw_joints2D=raw['joints2D'], raw_zrot=raw['zrot'],
                    poses=poses[::skip], shapes=shapes[::skip], joints2d=j2d[::skip].astype(np.float32), zrot=zrot[::skip])
print('ok', poses[::skip].shape)
