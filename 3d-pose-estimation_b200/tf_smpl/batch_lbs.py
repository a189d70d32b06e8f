"""Drop-in for deps/tf_smpl/batch_lbs.py (batch_skew, batch_rodrigues, batch_lrotmin,
batch_global_rigid_transformation) on the sm_100a kernels."""
import torch

from smplb200 import api


def batch_skew(vec, batch_size=None):
    """vec N x 3 -> N x 3 x 3 skew-symmetric matrices (batch_lbs.py:11-35).  Pure index shuffling
    (no arithmetic), done with carrier ops."""
    vec = torch.as_tensor(vec, dtype=torch.float32)
    z = torch.zeros_like(vec[:, 0])
    return torch.stack([z, -vec[:, 2], vec[:, 1], vec[:, 2], z, -vec[:, 0],
                        -vec[:, 1], vec[:, 0], z], dim=1).reshape(-1, 3, 3)


def batch_rodrigues(theta, name=None):
    """theta N x 3 -> N x 3 x 3 (batch_lbs.py:38-57)."""
    return api.rodrigues(theta, 0)


def batch_lrotmin(theta, name=None):
    """theta N x 72 -> N x 207: rotation matrices of the 23 non-root joints minus identity (batch_lbs.py:60-81)."""
    theta = torch.as_tensor(theta, dtype=torch.float32)
    Rs = api.rodrigues(theta[:, 3:].reshape(-1, 3).contiguous(), 0)
    return (Rs - torch.eye(3, dtype=Rs.dtype, device=Rs.device)).reshape(-1, 207)


def batch_global_rigid_transformation(Rs, Js, parent, rotate_base=False):
    """Rs N x 24 x 3 x 3, Js N x 24 x 3 -> (new_J N x 24 x 3, A N x 24 x 4 x 4) (batch_lbs.py:84-144)."""
    return api.global_rigid(Rs, Js, parent, rotate_base)
