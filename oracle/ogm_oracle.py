"""CPU oracle for occupancy-grid mapping -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline legs may import
this module; the product (``lidar_slam_b200``) never does.

Numpy restatement of the reference's OccupancyGridMap (modules/ogm.py):

* bresenham2D ........ modules/ogm.py:197-231 (same as code/pr2_utils.py:46-85)
* world2grid ......... modules/ogm.py:104-126
* update_map ......... modules/ogm.py:149-195
* build_map .......... modules/ogm.py:46-65

Parity pinning: (1) the reference's own known-answer vectors for Bresenham
(code/pr2_utils.py:94 and :96) are replayed by tests/test_ogm_oracle.py; (2) whole-map outputs
of the unmodified modules/ogm.py on seeded inputs are committed as tests/golden/ogm_golden.npz
by tests/golden/make_ogm_golden.py and replayed bit for bit.
"""
from __future__ import annotations

import numpy as np

__all__ = ["bresenham2D", "OgmOracle"]

LOG_ODDS = np.log(4.)                      # ogm.py:44
P_RL_XY = np.array([0.13323, 0.])          # ogm.py:165 (first two components)


def bresenham2D(sx, sy, ex, ey):
    """(n, 2) integer cells of the line, start to end inclusive (ogm.py:197-231)."""
    sx, sy, ex, ey = int(sx), int(sy), int(ex), int(ey)
    dx, dy = abs(ex - sx), abs(ey - sy)
    steep = dy > dx
    if steep:
        dx, dy = dy, dx
    if dy == 0:
        q = np.zeros(dx + 1, dtype=np.int64)
    else:
        # error term e_k = (floor(dx/2) - k*dy) mod dx; a step of the minor axis happens where
        # it wraps around, i.e. where the successive difference is >= 0           (ogm.py:214)
        e = np.mod(dx // 2 - dy * np.arange(dx + 1, dtype=np.int64), dx)
        q = np.append(0, (np.diff(e) >= 0).astype(np.int64))
    major = np.arange(dx + 1, dtype=np.int64)
    minor = np.cumsum(q)
    if steep:
        y = sy + major if sy <= ey else sy - major
        x = sx + minor if sx <= ex else sx - minor
    else:
        x = sx + major if sx <= ex else sx - major
        y = sy + minor if sy <= ey else sy - minor
    return np.stack((x, y), axis=1)


class OgmOracle:
    def __init__(self, resolution, max_x, max_y, min_x, min_y, buffer=1.):
        self.res, self.min_x, self.min_y = float(resolution), float(min_x), float(min_y)
        self.width = int(np.ceil((max_x - min_x) / resolution + buffer))       # ogm.py:37
        self.height = int(np.ceil((max_y - min_y) / resolution + buffer))      # ogm.py:38
        self.log_odds = np.zeros((self.width, self.height), dtype=np.float32)  # ogm.py:41
        self.grid = np.zeros((self.width, self.height), dtype=np.uint8)

    def world2grid(self, x, y):
        gx = np.ceil((np.asarray(x) - self.min_x) / self.res).astype(np.int32) - 1   # ogm.py:118
        gy = np.ceil((np.asarray(y) - self.min_y) / self.res).astype(np.int32) - 1
        return gx, gy

    def update_map(self, x_t, z_t):
        x, y, yaw = x_t
        R = np.array([[np.cos(yaw), -np.sin(yaw)], [np.sin(yaw), np.cos(yaw)]])      # ogm.py:159
        z_w = np.dot(z_t, R.T) + np.array([x, y])                                    # ogm.py:160
        s = np.array([x, y]) + P_RL_XY                                               # ogm.py:166 (offset not rotated)
        sx, sy = self.world2grid(s[0], s[1])
        ex, ey = self.world2grid(z_w[:, 0], z_w[:, 1])
        for i in range(z_w.shape[0]):
            pts = bresenham2D(sx, sy, ex[i], ey[i])
            ok = (pts[:, 0] >= 0) & (pts[:, 0] < self.width) & (pts[:, 1] >= 0) & (pts[:, 1] < self.height)
            pts = pts[ok]
            if pts.shape[0] == 0:
                continue
            # float32 array, float64 scalar: every update is f32(f64(cell) -/+ log 4)   (ogm.py:186-187)
            self.log_odds[pts[:-1, 0], pts[:-1, 1]] -= LOG_ODDS
            self.log_odds[pts[-1, 0], pts[-1, 1]] += LOG_ODDS
        self.log_odds = np.clip(self.log_odds, a_min=-20, a_max=20)                  # ogm.py:191

    def build_map(self, states, meas):
        for x_t, z_t in zip(states, meas):
            self.update_map(x_t, z_t)
        with np.errstate(over="ignore"):
            pmf = 1. / (1 + np.exp(self.log_odds))                                   # ogm.py:62
        self.grid[pmf > 0.5] = 1                                                     # ogm.py:63-65 (0.5 -> uint8 0)
        self.grid[pmf <= 0.5] = 0
