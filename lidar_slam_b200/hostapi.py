"""numpy-in / numpy-out binding of ``b200icp_run_host`` -- no torch anywhere on this path.

This is the binding a maintainer of the reference would write: the reference keeps its scans as a
list of ragged ``(n_i, 2)`` float64 arrays (modules/localization.py:252-307) and calls
``run_icp`` pair after pair (modules/localization.py:112-130, main.py:161-172).  Here the scans
are packed once into a padded pool and the whole pair list is handed to the library, which owns
the device memory, the host<->device copies and their overlap with the kernels.
"""
from __future__ import annotations

import ctypes
import weakref
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


def pinned_empty(shape, dtype) -> np.ndarray:
    """np.empty in page-locked memory (b200icp_host_alloc): the library's copies from and to it are real
    asynchronous DMA instead of driver-staged, blocking copies.  Freed when the array is collected.
    Falls back to pageable memory when the allocation fails (the calls still work, only slower)."""
    shape = tuple(int(x) for x in (shape if isinstance(shape, (tuple, list)) else (shape,)))
    dt = np.dtype(dtype)
    nbytes = max(int(np.prod(shape)) * dt.itemsize, 1)
    lib = _lib.load()
    ptr = lib.b200icp_host_alloc(nbytes)
    if not ptr:
        return np.empty(shape, dtype=dt)
    raw = (ctypes.c_char * nbytes).from_address(ptr)
    arr = np.frombuffer(raw, dtype=dt, count=int(np.prod(shape))).reshape(shape)
    weakref.finalize(raw, lib.b200icp_host_free, ctypes.c_void_p(ptr))
    return arr


def _p(a: Optional[np.ndarray]):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


def run_icp_batch_host(scans: np.ndarray, counts: np.ndarray, pairs: np.ndarray, init_transforms=None,
                       epsilon=0.01, max_iters: int = 2000, stopping_thresh: float = 1e-4,
                       normalize_error=False, return_correspondences: bool = False, flags: int = 0,
                       peer_pose_ptrs=None, peer_row0: int = 0, peer_flag=None) -> HostBatchResult:
    """run_icp (modules/icp.py:123-189) for every (source, target) row of `pairs` over a padded scan
    pool in host memory; same per-pair semantics and defaults as the reference call.
    flags: B200ICP_FLAG_* (_lib.FLAG_*); peer_pose_ptrs / peer_row0: the fused multi-GPU pose exchange
    (device pointers of every rank's gather buffer, include/b200icp.h); peer_flag: optional
    (flag_offset_bytes, flag_index, flag_value) arrival flags."""
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
    ex, keep = _lib.make_exchange(peer_pose_ptrs, peer_row0, *(peer_flag or (0, 0, 0)))
    _lib.check(lib.b200icp_run_host_ex(_p(scans), _p(counts), n_scans, n_max, dim, _p(pairs), _p(init), _p(eps),
                                       _p(norm), n_pairs, int(max_iters), float(stopping_thresh), _p(T), _p(err),
                                       _p(iters), _p(status), _p(corr), None if ex is None else ctypes.byref(ex),
                                       int(flags)), "b200icp_run_host_ex")
    return HostBatchResult(T, err, iters, status, corr)


def run_icp_pair_host(src: np.ndarray, tgt: np.ndarray, init: np.ndarray, epsilon: float, max_iters: int,
                      stopping_thresh: float, normalize: bool, want_corr: bool):
    """ONE pair through b200icp_run_pair_host (the C form of the reference's run_icp call): contiguous
    float64 (n, dim) clouds of one dimension, init (4, 4).  Returns (T, error, passes, status, corr|None)."""
    lib = _lib.load()
    T = np.empty((4, 4), dtype=np.float64)
    err = np.empty((1,), dtype=np.float64)
    meta = np.empty((2,), dtype=np.int32)
    corr = np.empty((src.shape[0],), dtype=np.int32) if want_corr else None
    _lib.check(lib.b200icp_run_pair_host(_p(src), src.shape[0], _p(tgt), tgt.shape[0], src.shape[1], _p(init),
                                         float(epsilon), int(max_iters), float(stopping_thresh), 1 if normalize else 0,
                                         _p(T), _p(err), ctypes.c_void_p(meta.ctypes.data),
                                         ctypes.c_void_p(meta.ctypes.data + 4), _p(corr)), "b200icp_run_pair_host")
    return T, err[0], int(meta[0]), int(meta[1]), corr


def beam_tables(n_beams: int):
    """cos / sin of linspace(-135 deg, +135 deg, n_beams) exactly as the reference evaluates them
    (modules/localization.py:273-278), numpy on the host."""
    angles = np.linspace(-135 * (np.pi / 180), 135 * (np.pi / 180), n_beams)
    return np.cos(angles), np.sin(angles)


class HostRangesWorkspace:
    """Reusable, page-locked host-side buffers for run_icp_ranges_host: results land in the same arrays on
    every call, so that a caller who loops pays no allocation and the copies back are asynchronous DMA;
    scalar epsilon / normalize_error are expanded into pinned per-pair arrays kept here.  Pass `out=`."""

    def __init__(self, n_pairs: int, n_scans: int, n_max: int, correspondences: bool = False):
        self.T = pinned_empty((n_pairs, 4, 4), np.float64)
        self.err = pinned_empty((n_pairs,), np.float64)
        self.iters = pinned_empty((n_pairs,), np.int32)
        self.status = pinned_empty((n_pairs,), np.int32)
        self.corr = pinned_empty((n_pairs, n_max), np.int32) if correspondences else None
        self.counts = pinned_empty((n_scans,), np.int32)
        self.eps = pinned_empty((n_pairs,), np.float64)
        self.norm = pinned_empty((n_pairs,), np.uint8)
        self._tables = {}

    def beam_tables(self, n_beams: int):
        """beam_tables(n_beams) in pinned memory, computed once."""
        if n_beams not in self._tables:
            c, s_ = beam_tables(n_beams)
            pc, ps = pinned_empty((n_beams,), np.float64), pinned_empty((n_beams,), np.float64)
            pc[:], ps[:] = c, s_
            self._tables[n_beams] = (pc, ps)
        return self._tables[n_beams]

    def per_pair(self, epsilon, normalize_error):
        """(eps, norm) per-pair arrays: the pinned ones filled with a scalar, or the caller's arrays."""
        n = self.eps.shape[0]
        eps = np.asarray(epsilon, dtype=np.float64)
        if eps.ndim == 0:
            self.eps.fill(float(eps)); eps = self.eps
        norm = np.asarray(normalize_error, dtype=np.uint8)
        if norm.ndim == 0:
            self.norm.fill(int(norm)); norm = self.norm
        return _per_pair(eps, n, np.float64), _per_pair(norm, n, np.uint8)


def run_icp_ranges_host(ranges: np.ndarray, pairs: np.ndarray, init_transforms=None, epsilon=0.01,
                        max_iters: int = 2000, stopping_thresh: float = 1e-4, normalize_error=False,
                        return_correspondences: bool = False, range_min: float = 0.1, range_max: float = 30.0,
                        offset_xy=(0.13323, 0.0), flags: int = 0, peer_pose_ptrs=None, peer_row0: int = 0,
                        peer_flag=None, out: Optional[HostRangesWorkspace] = None, tables=None):
    """get_lidar_data (modules/localization.py:252-307) + run_icp (modules/icp.py:123-189) for every
    (source, target) row of `pairs`, from the synced lidar ranges [S, n_beams] in host memory (float64
    as the reference holds them, or a float32 copy): b200icp_run_host_ranges.  Returns
    (HostBatchResult, counts [S] int32)."""
    lib = _lib.load()
    if ranges.dtype not in (np.float64, np.float32):
        ranges = np.asarray(ranges, dtype=np.float64)
    ranges = np.ascontiguousarray(ranges)
    if ranges.ndim != 2 or ranges.shape[1] == 0 or ranges.shape[1] > 2048:
        raise ValueError("ranges must be [S, n_beams] with n_beams <= 2048")
    n_scans, n_beams = ranges.shape
    n_max = n_beams + (n_beams & 1)
    pairs = np.ascontiguousarray(pairs, dtype=np.int32).reshape(-1, 2)
    n_pairs = pairs.shape[0]
    o = out if out is not None else HostRangesWorkspace(n_pairs, n_scans, n_max, return_correspondences)
    cos_t, sin_t = tables if tables is not None else o.beam_tables(n_beams)
    cos_t, sin_t = np.ascontiguousarray(cos_t, dtype=np.float64), np.ascontiguousarray(sin_t, dtype=np.float64)
    if init_transforms is None:
        init = np.broadcast_to(np.eye(4), (n_pairs, 4, 4)).copy()
    else:
        init = np.ascontiguousarray(init_transforms, dtype=np.float64).reshape(-1, 4, 4)
        if init.shape[0] != n_pairs:
            raise ValueError("init_transforms must have one 4x4 per pair")
    if o.T.shape[0] != n_pairs or o.counts.shape[0] != n_scans or (return_correspondences and o.corr is None):
        raise ValueError("`out` was built for another batch shape")
    eps, norm = o.per_pair(epsilon, normalize_error)
    ex, keep = _lib.make_exchange(peer_pose_ptrs, peer_row0, *(peer_flag or (0, 0, 0)))
    _lib.check(lib.b200icp_run_host_ranges(
        _p(ranges), 1 if ranges.dtype == np.float32 else 0, n_scans, n_beams, _p(cos_t), _p(sin_t), float(range_min),
        float(range_max), float(offset_xy[0]), float(offset_xy[1]), _p(pairs), _p(init), _p(eps), _p(norm), n_pairs,
        int(max_iters), float(stopping_thresh), _p(o.T), _p(o.err), _p(o.iters), _p(o.status),
        _p(o.corr) if return_correspondences else None, _p(o.counts),
        None if ex is None else ctypes.byref(ex), int(flags)), "b200icp_run_host_ranges")
    return HostBatchResult(o.T, o.err, o.iters, o.status, o.corr if return_correspondences else None), o.counts


class HostTicket:
    """A submitted run_icp_ranges_host call (b200icp_submit_host_ranges): `wait()` returns what the
    synchronous call returns.  Holds every array the library still reads or writes."""

    def __init__(self, ticket: int, out: HostRangesWorkspace, want_corr: bool, keep):
        self._ticket, self._out, self._want_corr, self._keep = ticket, out, want_corr, keep

    def wait(self):
        if self._ticket is not None:
            t, self._ticket = self._ticket, None
            _lib.check(_lib.load().b200icp_wait_host(t), "b200icp_wait_host")
            self._keep = None
        o = self._out
        return HostBatchResult(o.T, o.err, o.iters, o.status, o.corr if self._want_corr else None), o.counts


def submit_icp_ranges_host(ranges: np.ndarray, pairs: np.ndarray, init_transforms, out: HostRangesWorkspace,
                           epsilon=0.01, max_iters: int = 2000, stopping_thresh: float = 1e-4, normalize_error=False,
                           return_correspondences: bool = False, range_min: float = 0.1, range_max: float = 30.0,
                           offset_xy=(0.13323, 0.0), flags: int = 0, peer_pose_ptrs=None, peer_row0: int = 0,
                           peer_flag=None) -> HostTicket:
    """run_icp_ranges_host without the wait: returns a HostTicket at once; up to two calls may be in flight
    per device, each with a HostRangesWorkspace of its own (`out`, required: the results land there).  The
    upload and the first pairs of the next call overlap the queue tail and the result copies of this one.
    `ranges`, `pairs` and `init_transforms` must be C-contiguous arrays of the right dtype (float64 / float32,
    int32, float64) -- ideally page-locked (pinned_empty) -- and must not be modified before `wait()`."""
    lib = _lib.load()
    if ranges.dtype not in (np.float64, np.float32) or not ranges.flags.c_contiguous or ranges.ndim != 2 \
            or ranges.shape[1] == 0 or ranges.shape[1] > 2048:
        raise ValueError("ranges must be a C-contiguous float64/float32 [S, n_beams] array with n_beams <= 2048")
    n_scans, n_beams = ranges.shape
    n_max = n_beams + (n_beams & 1)
    if pairs.dtype != np.int32 or not pairs.flags.c_contiguous or pairs.ndim != 2 or pairs.shape[1] != 2:
        raise ValueError("pairs must be a C-contiguous int32 [B, 2] array")
    n_pairs = pairs.shape[0]
    init = init_transforms
    if init.dtype != np.float64 or not init.flags.c_contiguous or init.shape != (n_pairs, 4, 4):
        raise ValueError("init_transforms must be a C-contiguous float64 [B, 4, 4] array")
    if out.T.shape[0] != n_pairs or out.counts.shape[0] != n_scans or (return_correspondences and out.corr is None):
        raise ValueError("`out` was built for another batch shape")
    cos_t, sin_t = out.beam_tables(n_beams)
    eps, norm = out.per_pair(epsilon, normalize_error)
    ex, keep = _lib.make_exchange(peer_pose_ptrs, peer_row0, *(peer_flag or (0, 0, 0)))
    ticket = ctypes.c_int(-1)
    _lib.check(lib.b200icp_submit_host_ranges(
        _p(ranges), 1 if ranges.dtype == np.float32 else 0, n_scans, n_beams, _p(cos_t), _p(sin_t), float(range_min),
        float(range_max), float(offset_xy[0]), float(offset_xy[1]), _p(pairs), _p(init), _p(eps), _p(norm), n_pairs,
        int(max_iters), float(stopping_thresh), _p(out.T), _p(out.err), _p(out.iters), _p(out.status),
        _p(out.corr) if return_correspondences else None, _p(out.counts),
        None if ex is None else ctypes.byref(ex), int(flags), ctypes.byref(ticket)), "b200icp_submit_host_ranges")
    return HostTicket(ticket.value, out, return_correspondences, (ranges, pairs, init, eps, norm, cos_t, sin_t, ex, keep))


def release() -> None:
    """Free the library's device arena on the current device."""
    _lib.check(_lib.load().b200icp_host_release(), "b200icp_host_release")
