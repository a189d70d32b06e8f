"""Drop-in for deps/tf_smpl/projection.py."""
from smplb200 import api


def batch_orth_proj_idrot(X, camera, name=None):
    """X N x P x 3, camera N x 3 -> N x P x 2 (projection.py:10-26)."""
    return api.orth_proj(X, camera)
