"""Stub: deps/pose_3d/utils.py imports tf_pose.common but never uses it (only in comments, utils.py:239-240)."""
