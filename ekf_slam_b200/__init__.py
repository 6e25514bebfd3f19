"""ekf_slam_b200 -- B200-native EKF-SLAM filter core behind the reference's class interface.

``EKF_SLAM``, ``EKF_SLAM_UC``, ``Correspondence``, ``Landmark`` mirror SamShue/EKF_SLAM's MATLAB
classes; the arithmetic runs in hand-written sm_100a CUDA kernels (libekfslam_b200.so, C ABI in
include/ekfslam_b200.h).  Importing this package needs the built library; there is no CPU fallback.
"""
from ._lib import load as load_library, EkfError, LIB_PATH            # noqa: F401
from .filters import EKF_SLAM, EKF_SLAM_UC, Correspondence, Landmark  # noqa: F401
from .batch import FilterBatch, shard_filters                         # noqa: F401
from . import synth, shard, batch                                     # noqa: F401
