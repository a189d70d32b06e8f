import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, '3d-pose-estimation_b200')
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: needs a CUDA device (run on the B200 box)')


@pytest.fixture(scope='session')
def synth_dd():
    from smplb200 import synthetic as syn
    return syn.make_synthetic_smpl(0)


@pytest.fixture(scope='session')
def oracle64(synth_dd):
    from oracle import ref_numpy as rn
    return rn.load_model(synth_dd, dtype=np.float64)


@pytest.fixture(scope='session')
def oracle32(synth_dd):
    from oracle import ref_numpy as rn
    return rn.load_model(synth_dd, dtype=np.float32)


@pytest.fixture(scope='session')
def gpu_model(synth_dd):
    import torch
    if not torch.cuda.is_available():
        pytest.skip('no CUDA device')
    from smplb200 import Model
    return Model(synth_dd, device=0)


def maxrel(a, b):
    """max|a-b| / max|b| — the tolerance definition of SURVEY.md §8c."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))
