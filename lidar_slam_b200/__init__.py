"""lidar_slam_b200 -- B200-native batched ICP scan matching, a drop-in for the ICP hot
path of fadli0029/LiDAR-SLAM (modules/icp.py).  See DESIGN.md / INTEGRATION.md.

Names are resolved lazily so that ``lidar_slam_b200.hostapi`` (numpy in, numpy out over the C ABI's
host entry) can be used without importing torch; everything else uses torch for device memory.
"""
import importlib

__version__ = "0.1.0"

_LAZY = {
    "IcpBatchResult": "engine", "STATUS_NAMES": "engine", "nearest_batch": "engine", "pack_scans": "engine",
    "run_icp_batch": "engine", "voxel_downsample_device": "engine",
    "run_icp": "icp", "voxel_downsample": "icp",
    "run_icp_batch_host": "hostapi", "pack_scans_numpy": "hostapi",
}
__all__ = sorted(_LAZY)


def __getattr__(name):
    if name in _LAZY:
        value = getattr(importlib.import_module("." + _LAZY[name], __name__), name)
        globals()[name] = value
        return value
    raise AttributeError(f"module {__name__!r} has no attribute {name!r}")
