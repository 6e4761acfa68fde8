"""aslam-b200: B200-native EKF/UKF-SLAM filter core (hot path of iamarkaj/AwesomeSLAM).

The arithmetic lives in awesomeslam_b200/lib/libaslam_b200.so (hand-written sm_100a CUDA behind
the C ABI of include/aslam_c.h). Importing this package does not need a GPU; creating a filter
does, and there is no CPU fallback.
"""
from . import capi  # noqa: F401
from . import export  # noqa: F401
from .filters import BatchedFilter, EKFSlam, UKFSlam, Odometry, LargeFilter, LandMarks  # noqa: F401
from .replay import (Replay, load_landmarks_yaml, map_cfg1, map_cfg3, large_map_state, dense_P,  # noqa: F401
                     large_observations, laser_scans, laser_poses)

__version__ = "0.1.0"
