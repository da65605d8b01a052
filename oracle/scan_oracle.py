"""CPU oracle for scan construction -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline legs may import
this module; the product (``lidar_slam_b200``) never does.

Float64 numpy restatement of the two producers of the scan list the ICP path consumes:

* get_lidar_data ................ modules/localization.py:252-307
* statistical_outliers_removal .. modules/localization.py:223-250

Parity pinning: the reference holds no golden vectors for these functions, so the oracle is
pinned against outputs of the reference itself (``tests/golden/make_scan_golden.py`` imports
the unmodified ``/root/reference/modules/localization.py`` with a matplotlib stub and commits
inputs + outputs as ``tests/golden/scan_golden.npz``; ``tests/test_scan_oracle.py`` replays
them through this file bit for bit).
"""
from __future__ import annotations

import numpy as np

__all__ = ["get_lidar_data", "statistical_outliers_removal", "range_statistics"]

_P_RL_XY = np.array([0.13323, 0.0])   # localization.py:297 (z dropped at :304)


def get_lidar_data(lidar_ranges, lidar_range_min, lidar_range_max):
    """Ragged list of (n_i, 2) robot-frame end points, one per scan (localization.py:252-307)."""
    lidar_ranges = np.asarray(lidar_ranges)
    angle_min = -135 * (np.pi / 180)                       # :273
    angle_max = 135 * (np.pi / 180)                        # :274
    n_scans, n_meas = lidar_ranges.shape
    angles = np.linspace(angle_min, angle_max, n_meas)     # :278
    cos_a, sin_a = np.cos(angles), np.sin(angles)
    scans = []
    for i in range(n_scans):
        r = lidar_ranges[i]
        valid = (r >= lidar_range_min) & (r <= lidar_range_max)   # :288
        x = r[valid] * cos_a[valid]                               # :293
        y = r[valid] * sin_a[valid]                               # :294
        # :297-301: [x y 0] @ I + p_rl, i.e. one rounded addition per coordinate
        scans.append(np.column_stack((x + _P_RL_XY[0], y + _P_RL_XY[1])))
    return scans


def range_statistics(z_ts, k_std=3):
    """(mean, std, threshold) of the point ranges over all scans (localization.py:235-240)."""
    pts = np.vstack(z_ts)
    d = np.sqrt(np.sum(pts ** 2, axis=1))
    mean, std = np.mean(d), np.std(d)
    return mean, std, mean + k_std * std


def statistical_outliers_removal(z_ts, k_std=3):
    """Every scan keeps its points with range < mean + k_std * std (localization.py:223-250)."""
    _, _, thr = range_statistics(z_ts, k_std)
    out = []
    for scan in z_ts:
        d = np.sqrt(np.sum(scan ** 2, axis=1))             # :243
        out.append(scan[d < thr])                          # :244
    return out
