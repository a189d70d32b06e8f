"""Drop-in for deps/tf_perspective_projection/project.py on the sm_100a kernels."""
import torch

from smplb200 import api


class _Project(torch.autograd.Function):
    @staticmethod
    def forward(ctx, pts, cpos, cang, foc):
        ctx.save_for_backward(pts, cpos, cang, foc)
        return api.project(pts, cpos, cang, foc, P=1, cam_tiled=True)

    @staticmethod
    def backward(ctx, g_uv):
        pts, cpos, cang, foc = ctx.saved_tensors
        # the reference project() is differentiable in all four arguments (project.py:22-46)
        g_pts, g_cp, g_ca, g_f = api.project_backward(pts, cpos, cang, foc, g_uv.contiguous(), P=1, cam_tiled=True,
                                                      want_focal=True)
        return g_pts, g_cp, g_ca, g_f


def project(points_world_space, cams_pos, cams_angles, focal_lens):
    """points (batch, 3) + per-point camera -> (batch, 2) (project.py:7-48)."""
    dev = points_world_space.device if torch.is_tensor(points_world_space) and points_world_space.is_cuda \
        else torch.device('cuda', torch.cuda.current_device())

    def t(x):
        return torch.as_tensor(x, dtype=torch.float32).to(dev).contiguous()

    return _Project.apply(t(points_world_space), t(cams_pos), t(cams_angles), t(focal_lens))


def rodrigues_batch(rvecs):
    """(batch, 3) axis-angle -> (batch, 3, 3) (project.py:51-81)."""
    return api.rodrigues(rvecs, 1)
