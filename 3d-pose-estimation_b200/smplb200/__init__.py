"""smplb200 — B200-native (sm_100a) batched SMPL forward/backward + projection.

Python here is glue; the arithmetic lives in libsmplb200.so (C ABI: include/smplb200.h).
"""
from . import _lib  # noqa: F401
from .api import (Model, rodrigues, global_rigid, project, project_backward, orth_proj,  # noqa: F401
                  reproj_loss, adam_step, mse_grad, add_axis_angle_rotations, rotate_global_pose, cam_losses)
from ._lib import F_BLEND_SIMT, F_KEEP_VPOSED, F_NO_SKIN, F_SKIN_SIMT, F_SCRATCH_VERTS, F_FUSED, SmplbError  # noqa: F401
