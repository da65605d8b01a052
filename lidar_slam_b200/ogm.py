"""Occupancy-grid mapping on the device.

``OccupancyGridMap`` mirrors the reference class of the same name (modules/ogm.py:5-231):
same constructor arguments, ``build_map(states, meas)``, ``update_map(x_t, z_t)``,
``world2grid`` / ``grid2world`` and the two result arrays ``grid_map`` (uint8) and
``grid_map_log_odds`` (float32) -- computed by ``b200icp_ogm_build``.  The matplotlib plotting
helpers (ogm.py:67-102) are out of scope.  ``build_maps`` is the batched entry the reference
lacks: several pose sequences (odometry / scan matching / optimised) over the same scans, one
map each, in one launch.
"""
from __future__ import annotations

import ctypes
from typing import Sequence

import numpy as np
import torch

from . import _lib
from .engine import _device, _ptr, _stream_ptr, pack_scans

_OCCUPIED_BELOW = None


def _reference_pmf_is_occupied(lo: np.ndarray) -> np.ndarray:
    """The reference's float32 expression (ogm.py:62-63) evaluated by numpy on the host."""
    with np.errstate(over="ignore"):
        return (1. / (1 + np.exp(lo.astype(np.float32)))) > 0.5


def one_at_or_below() -> np.float32:
    """Largest float32 log-odds value the reference maps to grid = 1 (pmf > 0.5, i.e. cells seen free).  The expression is
    monotone in the log-odds, so one bisection over the float32 bit patterns near zero finds
    the threshold the binarisation kernel compares against."""
    global _OCCUPIED_BELOW
    if _OCCUPIED_BELOW is None:
        lo_bits = np.array([np.float32(-1.0)]).view(np.uint32)[0]      # occupied
        hi_bits = np.array([np.float32(-1e-30)]).view(np.uint32)[0]    # not occupied (pmf == 0.5)
        assert _reference_pmf_is_occupied(np.array([-1.0]))[0]
        assert not _reference_pmf_is_occupied(np.array([-1e-30]))[0]
        # negative floats: larger magnitude = larger bit pattern; occupied for bits >= threshold
        a, b = int(hi_bits), int(lo_bits)                              # a: not occupied, b: occupied
        while b - a > 1:
            mid = (a + b) // 2
            v = np.array([mid], dtype=np.uint32).view(np.float32)
            if _reference_pmf_is_occupied(v)[0]:
                b = mid
            else:
                a = mid
        _OCCUPIED_BELOW = np.array([b], dtype=np.uint32).view(np.float32)[0]
    return _OCCUPIED_BELOW


def _as_pool(meas, dev):
    if isinstance(meas, (tuple, list)) and len(meas) == 2 and isinstance(meas[0], torch.Tensor) \
            and meas[0].dim() == 3:
        pool, counts = meas
        return pool.to(dev, torch.float64).contiguous(), counts.to(dev, torch.int32).contiguous()
    clouds = [np.asarray(z, dtype=np.float64).reshape(-1, 2) for z in meas]
    if max(c.shape[0] for c in clouds) == 0:
        clouds = [np.zeros((0, 2))] * len(clouds)
        pool = torch.zeros((len(clouds), 2, 2), dtype=torch.float64, device=dev)
        return pool, torch.zeros((len(clouds),), dtype=torch.int32, device=dev)
    return pack_scans(clouds, device=dev)


def build_maps(log_odds: torch.Tensor, res, min_x, min_y, poses, meas):
    """Update `log_odds` [M, W, H] float32 (device, in place) with every scan of `meas` seen from
    the M pose sequences `poses` [M, N, 3] (x, y, yaw).  `meas`: ragged list of (n_i, 2) arrays or
    a (pool, counts) pair on the device."""
    lib = _lib.load()
    dev = log_odds.device
    if not (log_odds.is_cuda and log_odds.dtype == torch.float32 and log_odds.is_contiguous() and log_odds.dim() == 3):
        raise ValueError("log_odds must be a contiguous CUDA float32 tensor [M, W, H]")
    n_maps, width, height = log_odds.shape
    poses_h = np.ascontiguousarray(np.asarray(poses, dtype=np.float64)).reshape(n_maps, -1, 3)
    n_scans = poses_h.shape[1]
    pool, counts = _as_pool(meas, dev)
    if pool.shape[0] != n_scans:
        raise ValueError("one scan per pose is required")
    # cos / sin of the yaw as numpy evaluates them (ogm.py:159), so that the end points match
    cs_h = np.stack([np.cos(poses_h[..., 2]), np.sin(poses_h[..., 2])], axis=-1)
    poses_d = torch.as_tensor(poses_h).to(dev)
    cs_d = torch.as_tensor(np.ascontiguousarray(cs_h)).to(dev)
    n_max = pool.shape[1]
    nbytes = ctypes.c_size_t()
    _lib.check(lib.b200icp_ogm_workspace_bytes(width, height, n_max, n_maps, ctypes.byref(nbytes)),
               "b200icp_ogm_workspace_bytes")
    ws = torch.empty((nbytes.value,), dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        _lib.check(lib.b200icp_ogm_build(_ptr(log_odds), width, height, float(res), float(min_x), float(min_y),
                                         _ptr(poses_d), _ptr(cs_d), _ptr(pool), _ptr(counts), n_scans, n_max,
                                         n_maps, _ptr(ws), ws.numel(), _stream_ptr()), "b200icp_ogm_build")
    return log_odds


def binarize(log_odds: torch.Tensor) -> torch.Tensor:
    """grid_map of build_map (ogm.py:62-65): 1 where pmf > 0.5, else 0 (uint8)."""
    lib = _lib.load()
    grid = torch.empty(log_odds.shape, dtype=torch.uint8, device=log_odds.device)
    with torch.cuda.device(log_odds.device):
        _lib.check(lib.b200icp_ogm_binarize(_ptr(log_odds), _ptr(grid), log_odds.numel(),
                                            ctypes.c_float(float(one_at_or_below())), _stream_ptr()),
                   "b200icp_ogm_binarize")
    return grid


class OccupancyGridMap:
    """Same public surface as the reference's OccupancyGridMap (modules/ogm.py:5-231)."""

    def __init__(self, resolution, world_map_max_x, world_map_max_y, world_map_min_x, world_map_min_y,
                 buffer=1., device=None):
        self.res = float(resolution)
        self.world_map_max_x = float(world_map_max_x)
        self.world_map_max_y = float(world_map_max_y)
        self.world_map_min_x = float(world_map_min_x)
        self.world_map_min_y = float(world_map_min_y)
        self.buffer = float(buffer)
        # ogm.py:37-38
        self.grid_map_width = int(np.ceil((world_map_max_x - world_map_min_x) / resolution + buffer))
        self.grid_map_height = int(np.ceil((world_map_max_y - world_map_min_y) / resolution + buffer))
        self.logodds_ratio = np.log(4.)
        self._dev = torch.device(device) if device is not None else _device()
        self._log_odds = torch.zeros((1, self.grid_map_width, self.grid_map_height), dtype=torch.float32,
                                     device=self._dev)
        self._grid = torch.zeros((self.grid_map_width, self.grid_map_height), dtype=torch.uint8, device=self._dev)

    # results, as numpy like the reference's attributes
    @property
    def grid_map_log_odds(self) -> np.ndarray:
        return self._log_odds[0].cpu().numpy()

    @property
    def grid_map(self) -> np.ndarray:
        return self._grid.cpu().numpy()

    def build_map(self, states, meas):
        """ogm.py:46-65: all scans in order, then the occupancy thresholding."""
        states = np.asarray(states, dtype=np.float64).reshape(-1, 3)
        build_maps(self._log_odds, self.res, self.world_map_min_x, self.world_map_min_y, states[None], meas)
        self._grid = binarize(self._log_odds[0])

    def update_map(self, x_t, z_t):
        """ogm.py:149-195: one scan."""
        build_maps(self._log_odds, self.res, self.world_map_min_x, self.world_map_min_y,
                   np.asarray(x_t, dtype=np.float64).reshape(1, 1, 3), [np.asarray(z_t, dtype=np.float64)])

    def world2grid(self, x, y):
        """ogm.py:104-126 (host arithmetic, identical to the reference)."""
        if isinstance(x, np.ndarray) and isinstance(y, np.ndarray):
            x_g = np.ceil((x - self.world_map_min_x) / self.res).astype(np.int32) - 1
            y_g = np.ceil((y - self.world_map_min_y) / self.res).astype(np.int32) - 1
            return np.hstack((x_g.reshape(-1, 1), y_g.reshape(-1, 1)))
        i = int(np.ceil((x - self.world_map_min_x) / self.res)) - 1
        j = int(np.ceil((y - self.world_map_min_y) / self.res)) - 1
        return np.array([i, j], dtype=np.int32)

    def grid2world(self, i, j):
        """ogm.py:128-147."""
        x = i * self.res + self.world_map_min_x
        y = j * self.res + self.world_map_min_y
        if isinstance(i, np.ndarray) and isinstance(j, np.ndarray):
            return np.hstack((x.reshape(-1, 1), y.reshape(-1, 1)))
        return np.array([x, y])
