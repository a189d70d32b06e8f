"""Drop-in for the tensor helpers of deps/pose_3d (utils.py, config.py) on the sm_100a kernels.  The CNN
(network.py), the training loop and the tf.data pipeline of the reference are out of scope."""
from . import config, utils  # noqa: F401
