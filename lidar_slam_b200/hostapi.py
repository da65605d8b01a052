"""numpy-in / numpy-out binding of ``b200icp_run_host`` -- no torch anywhere on this path.

This is the binding a maintainer of the reference would write: the reference keeps its scans as a
list of ragged ``(n_i, 2)`` float64 arrays (modules/localization.py:252-307) and calls
``run_icp`` pair after pair (modules/localization.py:112-130, main.py:161-172).  Here the scans
are packed once into a padded pool and the whole pair list is handed to the library, which owns
the device memory, the host<->device copies and their overlap with the kernels.
"""
from __future__ import annotations

import ctypes
from typing import NamedTuple, Optional, Sequence

import numpy as np

from . import _lib


class HostBatchResult(NamedTuple):
    T: np.ndarray            # [B, 4, 4] float64
    error: np.ndarray        # [B] float64
    iters: np.ndarray        # [B] int32
    status: np.ndarray       # [B] int32 (engine.STATUS_NAMES)
    correspondences: Optional[np.ndarray]   # [B, n_max] int32 or None


def pack_scans_numpy(clouds: Sequence[np.ndarray]):
    """Ragged list of (n_i, D) arrays -> (pool [S, n_max, D] float64, counts [S] int32); n_max is
    rounded up to an even count (16-byte aligned rows), the layout engine.pack_scans produces."""
    clouds = [np.asarray(c, dtype=np.float64) for c in clouds]
    if not clouds:
        raise ValueError("at least one scan is required")
    dim = clouds[0].shape[1] if clouds[0].ndim == 2 else 0
    if dim not in (2, 3) or any(c.ndim != 2 or c.shape[1] != dim for c in clouds):
        raise ValueError("scans must be (n_i, 2) or (n_i, 3) arrays of one dimension")
    n_max = max(max(c.shape[0] for c in clouds), 1)
    n_max += n_max % 2
    pool = np.zeros((len(clouds), n_max, dim), dtype=np.float64)
    counts = np.zeros((len(clouds),), dtype=np.int32)
    for i, c in enumerate(clouds):
        pool[i, :c.shape[0]] = c
        counts[i] = c.shape[0]
    return pool, counts


def _per_pair(value, n, dtype):
    a = np.asarray(value, dtype=dtype)
    if a.ndim == 0:
        a = np.full((n,), a, dtype=dtype)
    if a.shape != (n,):
        raise ValueError("per-pair arguments must be scalars or have one entry per pair")
    return np.ascontiguousarray(a)


def _p(a: Optional[np.ndarray]):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


def run_icp_batch_host(scans: np.ndarray, counts: np.ndarray, pairs: np.ndarray, init_transforms=None,
                       epsilon=0.01, max_iters: int = 2000, stopping_thresh: float = 1e-4,
                       normalize_error=False, return_correspondences: bool = False) -> HostBatchResult:
    """run_icp (modules/icp.py:123-189) for every (source, target) row of `pairs` over a padded scan
    pool in host memory; same per-pair semantics and defaults as the reference call."""
    lib = _lib.load()
    scans = np.ascontiguousarray(scans, dtype=np.float64)
    if scans.ndim != 3 or scans.shape[2] not in (2, 3):
        raise ValueError("scans must be [S, n_max, 2|3] float64 (see pack_scans_numpy)")
    n_scans, n_max, dim = scans.shape
    if (n_max * dim) % 2:
        raise ValueError("n_max*dim must be even (16-byte aligned rows); use pack_scans_numpy")
    counts = np.ascontiguousarray(counts, dtype=np.int32)
    if counts.shape != (n_scans,):
        raise ValueError("counts must have one entry per scan")
    pairs = np.ascontiguousarray(pairs, dtype=np.int32).reshape(-1, 2)
    n_pairs = pairs.shape[0]
    if init_transforms is None:
        init = np.broadcast_to(np.eye(4), (n_pairs, 4, 4)).copy()
    else:
        init = np.ascontiguousarray(init_transforms, dtype=np.float64).reshape(-1, 4, 4)
        if init.shape[0] != n_pairs:
            raise ValueError("init_transforms must have one 4x4 per pair")
    eps = _per_pair(epsilon, n_pairs, np.float64)
    norm = _per_pair(normalize_error, n_pairs, np.uint8)
    T = np.empty((n_pairs, 4, 4), dtype=np.float64)
    err = np.empty((n_pairs,), dtype=np.float64)
    iters = np.empty((n_pairs,), dtype=np.int32)
    status = np.empty((n_pairs,), dtype=np.int32)
    corr = np.empty((n_pairs, n_max), dtype=np.int32) if return_correspondences else None
    _lib.check(lib.b200icp_run_host(_p(scans), _p(counts), n_scans, n_max, dim, _p(pairs), _p(init), _p(eps),
                                    _p(norm), n_pairs, int(max_iters), float(stopping_thresh), _p(T), _p(err),
                                    _p(iters), _p(status), _p(corr)), "b200icp_run_host")
    return HostBatchResult(T, err, iters, status, corr)


def release() -> None:
    """Free the library's device arena on the current device."""
    _lib.check(_lib.load().b200icp_host_release(), "b200icp_host_release")
