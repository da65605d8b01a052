"""Scan construction on the device: the producer of the scan pool.

Mirrors, with the same names, argument meaning and results,

* ``get_lidar_data``               modules/localization.py:252-307
* ``statistical_outliers_removal`` modules/localization.py:223-250

but keeps the scans in the padded device pool ``[S, n_max, 2]`` float64 + ``counts [S]`` that
``run_icp_batch`` reads, so the ragged host list -> padded tensor copy disappears.  The
list-of-arrays forms the reference returns are available through ``pool_to_list``.
DBSCAN (localization.py:201-221, sklearn) stays on the host and is out of scope.
"""
from __future__ import annotations

import ctypes
from typing import List, Tuple

import numpy as np
import torch

from . import _lib
from .engine import _device, _ptr, _stream_ptr

# localization.py:273-278 and :297 (the lidar sits 0.13323 m ahead of the robot centre)
ANGLE_MIN_DEG, ANGLE_MAX_DEG = -135.0, 135.0
LIDAR_OFFSET_XY = (0.13323, 0.0)


def beam_tables(n_beams: int) -> Tuple[np.ndarray, np.ndarray]:
    """cos / sin of the beam angles exactly as the reference evaluates them (numpy, host)."""
    angle_min = ANGLE_MIN_DEG * (np.pi / 180)
    angle_max = ANGLE_MAX_DEG * (np.pi / 180)
    angles = np.linspace(angle_min, angle_max, n_beams)
    return np.cos(angles), np.sin(angles)


def scans_from_ranges(lidar_ranges, lidar_range_min, lidar_range_max, device=None):
    """get_lidar_data on the device.

    lidar_ranges (N, n_beams) float (numpy or tensor, host or device) ->
    (pool [N, n_max, 2] float64, counts [N] int32) on the device; n_max = n_beams rounded up
    to an even count (16-byte aligned rows for the TMA bulk copies).
    """
    lib = _lib.load()
    dev = torch.device(device) if device is not None else (
        lidar_ranges.device if isinstance(lidar_ranges, torch.Tensor) and lidar_ranges.is_cuda else _device())
    if isinstance(lidar_ranges, torch.Tensor):
        r = lidar_ranges.to(device=dev, dtype=torch.float64, non_blocking=True).contiguous()
    else:
        r = torch.as_tensor(np.ascontiguousarray(lidar_ranges, dtype=np.float64)).to(dev)
    if r.dim() != 2 or r.shape[1] == 0:
        raise ValueError("lidar_ranges must have shape (N, n_beams)")
    n_scans, n_beams = r.shape
    n_max = n_beams + (n_beams & 1)
    c, s = beam_tables(n_beams)
    tab = torch.as_tensor(np.stack([c, s])).to(dev)
    pool = torch.empty((n_scans, n_max, 2), dtype=torch.float64, device=dev)
    counts = torch.empty((n_scans,), dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        _lib.check(lib.b200icp_scans_from_ranges(
            _ptr(r), n_scans, n_beams, _ptr(tab[0]), _ptr(tab[1]), float(lidar_range_min),
            float(lidar_range_max), LIDAR_OFFSET_XY[0], LIDAR_OFFSET_XY[1], _ptr(pool), _ptr(counts),
            n_max, _stream_ptr()), "b200icp_scans_from_ranges")
    return pool, counts


def statistical_outliers_removal_pool(pool: torch.Tensor, counts: torch.Tensor, k_std=3):
    """statistical_outliers_removal on the pool, in place.  Returns (pool, counts, stats) with
    stats = device tensor [mean, std, threshold, n_points] (what the reference prints)."""
    lib = _lib.load()
    if not (pool.is_cuda and pool.dtype == torch.float64 and pool.is_contiguous() and pool.dim() == 3
            and pool.shape[2] == 2):
        raise ValueError("pool must be a contiguous CUDA float64 tensor [S, n_max, 2]")
    if not (counts.is_cuda and counts.dtype == torch.int32 and counts.is_contiguous()):
        raise ValueError("counts must be a contiguous CUDA int32 tensor [S]")
    dev = pool.device
    nbytes = ctypes.c_size_t()
    _lib.check(lib.b200icp_range_filter_workspace_bytes(ctypes.byref(nbytes)), "b200icp_range_filter_workspace_bytes")
    ws = torch.empty((nbytes.value,), dtype=torch.uint8, device=dev)
    stats = torch.zeros((4,), dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        _lib.check(lib.b200icp_range_filter(_ptr(pool), _ptr(counts), pool.shape[0], pool.shape[1],
                                            float(k_std), _ptr(stats), _ptr(ws), ws.numel(),
                                            _stream_ptr()), "b200icp_range_filter")
    return pool, counts, stats


def pool_to_list(pool: torch.Tensor, counts: torch.Tensor) -> List[np.ndarray]:
    """Padded pool -> the reference's ragged list of (n_i, D) float64 arrays (host)."""
    p, c = pool.cpu().numpy(), counts.cpu().numpy()
    return [p[i, :c[i]].copy() for i in range(p.shape[0])]


def get_lidar_data(lidar_ranges, lidar_range_min, lidar_range_max) -> List[np.ndarray]:
    """Drop-in for modules/localization.py:252-307 (same signature, same list-of-arrays result)."""
    return pool_to_list(*scans_from_ranges(lidar_ranges, lidar_range_min, lidar_range_max))


def statistical_outliers_removal(z_ts, k_std=3) -> List[np.ndarray]:
    """Drop-in for modules/localization.py:223-250 on a ragged host list."""
    from .engine import pack_scans
    pool, counts = pack_scans(z_ts)
    pool, counts, _ = statistical_outliers_removal_pool(pool, counts, k_std)
    return pool_to_list(pool, counts)
