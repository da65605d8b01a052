"""lidar_slam_b200 -- B200-native batched ICP scan matching, a drop-in for the ICP hot
path of fadli0029/LiDAR-SLAM (modules/icp.py).  See DESIGN.md / INTEGRATION.md."""
from .engine import (IcpBatchResult, STATUS_NAMES, nearest_batch, pack_scans, run_icp_batch,  # noqa: F401
                     voxel_downsample_device)
from .icp import run_icp, voxel_downsample  # noqa: F401

__version__ = "0.1.0"
