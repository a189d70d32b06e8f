"""Constants of deps/pose_3d/config.py:6-24 that the tensor helpers and losses read."""
from math import pi

n_joints = 14
n_joints_smpl = 24
input_img_size = (240, 320)   # (height, width)
fl = 0.05
ss = 0.024
total_loss_scale = 1.0
pose_loss_scale = 1.0
pose_loss_direct_scale = 1.0
mesh_loss_scale = 1.0
joint_loss_scale = 2.0
reg_loss_scale = 100.0
joint_limit = pi * 2.0
disc_loss_scale = 2
reproj_loss_scale = 1.0 * ss
cam_loss_scale = 1 / (400 * n_joints_smpl)
cam_angle_loss_scale = 10.0
