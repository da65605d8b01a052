"""Host side of the batched ICP engine: device buffers (torch) + the C-ABI calls.

`run_icp_batch` is the new batched entry the reference lacks (its callers loop,
modules/localization.py:115-130, main.py:161-172, test_icp.py:70-91); per pair
it has exactly the semantics of the reference's run_icp (modules/icp.py:123-189).
PyTorch is used for device memory and streams only.
"""
from __future__ import annotations

import ctypes
import os
from dataclasses import dataclass
from typing import Optional, Sequence, Union

import numpy as np
import torch

from . import _lib

STATUS_NAMES = {
    0: "epsilon", 1: "max_iters", 2: "stalled", 3: "nonfinite", 4: "empty", 5: "nonplanar_init",
    6: "skipped", 7: "not_resident",
}
FLAG_EXHAUSTIVE, FLAG_HOST_CHUNKED, FLAG_NO_PRUNE = _lib.FLAG_EXHAUSTIVE, _lib.FLAG_HOST_CHUNKED, _lib.FLAG_NO_PRUNE
STATUS_NONPLANAR = 5


@dataclass
class IcpBatchResult:
    """Per-pair results, all on the device the inputs live on."""
    T: torch.Tensor          # [B,4,4] float64, transforms[-1] of icp.py:183-189
    error: torch.Tensor      # [B] float64, error of the last executed pass
    iters: torch.Tensor      # [B] int32, executed passes
    status: torch.Tensor     # [B] int32, see STATUS_NAMES
    corr: Optional[torch.Tensor]  # [B,n_max] int32 or None


def _device():
    if not torch.cuda.is_available():
        raise RuntimeError("lidar_slam_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback")
    return torch.device("cuda", torch.cuda.current_device())


def _ptr(t: Optional[torch.Tensor]):
    return ctypes.c_void_p(0 if t is None else t.data_ptr())


def _stream_ptr():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _to_dev(x, dtype, dev):
    """Host arrays / tensors -> contiguous device tensor (async when the source is pinned)."""
    if isinstance(x, torch.Tensor):
        return x.to(device=dev, dtype=dtype, non_blocking=True).contiguous()
    return torch.as_tensor(np.ascontiguousarray(x), dtype=dtype).to(dev)


def _keeps_plane(init: torch.Tensor) -> bool:
    """True if every 4x4 maps the z = 0 plane to itself with a unit homogeneous row, i.e. has the
    form TSE3_from_TSE2 produces (modules/utils.py:227-240)."""
    zero = init[:, [0, 1, 2, 2, 2, 3, 3, 3], [2, 2, 0, 1, 3, 0, 1, 2]]
    return bool(((zero == 0).all() & (init[:, 3, 3] == 1).all()).item())


def pack_scans(clouds: Sequence[np.ndarray], device=None):
    """Ragged list of (n_i, D) arrays -> padded pool [S, n_max, D] float64 + counts [S] int32.

    This is the layout of the scan pool in HBM: one row of n_max*D float64 per scan,
    valid points first.  n_max is rounded up to an even count so that every row
    starts 16-byte aligned for the TMA bulk copies.
    """
    device = device or _device()
    if len(clouds) == 0:
        raise ValueError("no point clouds given")
    dim = int(np.asarray(clouds[0]).shape[1])
    if dim not in (2, 3):
        raise ValueError("point clouds must have shape (N,2) or (N,3)")
    n_max = max(int(np.asarray(c).shape[0]) for c in clouds)
    if n_max == 0:
        raise ValueError("empty point cloud")
    n_max += n_max & 1
    pool = np.zeros((len(clouds), n_max, dim), dtype=np.float64)
    counts = np.zeros((len(clouds),), dtype=np.int32)
    for i, c in enumerate(clouds):
        c = np.asarray(c, dtype=np.float64)
        if c.ndim != 2 or c.shape[1] != dim:
            raise ValueError("all point clouds of a pool must share their dimension")
        pool[i, :c.shape[0]] = c
        counts[i] = c.shape[0]
    return torch.from_numpy(pool).to(device), torch.from_numpy(counts).to(device)


def _per_pair(x, n, dtype, dev):
    if isinstance(x, (int, float, bool, np.floating, np.integer, np.bool_)):
        return torch.full((n,), x, dtype=dtype, device=dev)
    t = _to_dev(x, dtype, dev).reshape(-1)
    if t.numel() != n:
        raise ValueError("per-pair parameter has the wrong length")
    return t


_PIPELINE_MIN_PAIRS = 2048
_PIPELINE_CHUNKS = 6
_SIDE_STREAMS = {}


def _side_streams(dev):
    key = str(dev)
    if key not in _SIDE_STREAMS:
        _SIDE_STREAMS[key] = [torch.cuda.Stream(device=dev) for _ in range(3)]     # copy, compute A, compute B
    return _SIDE_STREAMS[key]


_GATE_SLICES = 24
_GATE_VALUES = {}


def _copy_pool_gated(host_pool, dev_pool, dev, launch_gated):
    """Host -> device copy of the scan pool under ONE launch (b200icp_run_gated): the pool goes up in
    slices on the copy stream, each followed by a 4-byte copy that raises a device counter of resident
    scans; the kernel, launched right behind on the current stream, waits on that counter per pair."""
    cur = torch.cuda.current_stream(dev)
    copy_s = _side_streams(dev)[0]
    n_scans = host_pool.shape[0]
    k = max(1, min(_GATE_SLICES, n_scans))
    key = (n_scans, k)
    if key not in _GATE_VALUES:
        _GATE_VALUES[key] = torch.tensor([n_scans * (c + 1) // k for c in range(k)], dtype=torch.int32).pin_memory()
    vals = _GATE_VALUES[key]
    flag = torch.zeros((64,), dtype=torch.int32, device=dev)
    flag.record_stream(copy_s)
    copy_s.wait_stream(cur)            # counter zeroed, small inputs uploaded
    # copies are enqueued BEFORE the launch: a tool that serialises launches (profiler,
    # CUDA_LAUNCH_BLOCKING) must never leave the kernel waiting for a copy not yet issued
    with torch.cuda.stream(copy_s):
        lo = 0
        for c in range(k):
            hi = int(vals[c])
            if hi > lo:
                dev_pool[lo:hi].copy_(host_pool[lo:hi], non_blocking=True)
            flag[:1].copy_(vals[c:c + 1], non_blocking=True)
            lo = hi
    launch_gated(flag)
    cur.wait_stream(copy_s)


def _copy_pool_pipelined(host_pool, dev_pool, pairs_host, dev, launch_chunk):
    """Host -> device copy of the scan pool overlapped with the alignment: the pair list is cut
    into chunks; before chunk c runs, the prefix of the pool it references (rows up to its
    largest scan index) has been copied on the copy stream; chunks alternate between two
    compute streams so that one chunk's queue tail overlaps the next chunk's start.
    launch_chunk(lo, hi, workspace) launches pairs [lo, hi) on the current stream; None = copy only."""
    cur = torch.cuda.current_stream(dev)
    copy_s, comp_a, comp_b = _side_streams(dev)
    n_pairs, n_scans = pairs_host.shape[0], host_pool.shape[0]
    k = _PIPELINE_CHUNKS if launch_chunk is not None else 1
    bounds = [n_pairs * c // k for c in range(k + 1)]
    copy_s.wait_stream(cur)
    comp_a.wait_stream(cur)
    comp_b.wait_stream(cur)
    copied = 0
    for c in range(k):
        lo, hi = bounds[c], bounds[c + 1]
        need = n_scans if (launch_chunk is None or c == k - 1) else min(n_scans, int(pairs_host[lo:hi].max()) + 1)
        with torch.cuda.stream(copy_s):
            if need > copied:
                dev_pool[copied:need].copy_(host_pool[copied:need], non_blocking=True)
                copied = need
            ready = torch.cuda.Event()
            ready.record(copy_s)
        if launch_chunk is None or hi <= lo:
            continue
        st = comp_a if c % 2 == 0 else comp_b
        st.wait_event(ready)
        with torch.cuda.stream(st):
            nbytes = ctypes.c_size_t()
            _lib.check(_lib.load().b200icp_workspace_bytes(hi - lo, dev_pool.shape[1], dev_pool.shape[2],
                                                           ctypes.byref(nbytes)), "b200icp_workspace_bytes")
            ws_c = torch.empty((nbytes.value,), dtype=torch.uint8, device=dev)
            ws_c.record_stream(st)
            launch_chunk(lo, hi, ws_c)
    cur.wait_stream(copy_s)
    cur.wait_stream(comp_a)
    cur.wait_stream(comp_b)


def run_icp_batch(scans, counts, pairs, init_transforms=None, epsilon=0.01, max_iters=2000,
                  stopping_thresh=0.0001, normalize_error=False, return_correspondences=False,
                  planar_fallback=True, peer_pose_ptrs=None, peer_row0=0, flags=0, peer_flag=None) -> IcpBatchResult:
    """Align every (source, target) pair of `pairs` in one launch.

    scans  [S, n_max, D] float64 (D = 2 or 3), counts [S] int32 (see pack_scans);
    pairs  [B, 2] int32 rows (source scan, target scan);
    init_transforms [B,4,4] float64 (default identity);
    epsilon, normalize_error: scalar or length-B (so that odometry pairs and
    loop-closure pairs with different settings share one launch);
    max_iters, stopping_thresh: scalars, as modules/icp.py:128-129.
    peer_pose_ptrs: optional list of peer-mapped device pointers (one gather buffer per rank);
    the kernel then also stores every pair's (x, y, theta, error) row into row peer_row0 + b of
    each buffer (b200icp_run_exchange; see run_icp_sharded).
    peer_flag: optional (flag_offset_bytes, flag_index, flag_value): arrival flags of the exchange
    (b200icp_exchange in include/b200icp.h) instead of a cross-rank barrier after the launch.
    flags: B200ICP_FLAG_* (FLAG_EXHAUSTIVE, FLAG_NO_PRUNE: same results, for A/B timing and tests).

    Asynchronous on the current stream, with one exception: a big batch whose scan pool is a PINNED
    HOST tensor is streamed to the device under a gated launch, and that path synchronises before it
    returns (it must check that the kernel never gave up waiting for a slice), so the host pool may
    be refilled as soon as the call returns.
    """
    lib = _lib.load()
    dev = scans.device if isinstance(scans, torch.Tensor) and scans.is_cuda else _device()
    # Host inputs in pinned memory and a big batch: the scan pool is streamed to the device in
    # slices on a copy stream while earlier chunks of the pair list are already being aligned.
    host_pool = None
    if (isinstance(scans, torch.Tensor) and not scans.is_cuda and scans.is_pinned() and scans.dtype == torch.float64
            and scans.is_contiguous() and scans.dim() == 3 and not (isinstance(pairs, torch.Tensor) and pairs.is_cuda)):
        n_host_pairs = int(np.asarray(pairs).reshape(-1, 2).shape[0]) if not isinstance(pairs, torch.Tensor) \
            else int(pairs.reshape(-1, 2).shape[0])
        if n_host_pairs >= _PIPELINE_MIN_PAIRS:
            host_pool = scans
            scans = torch.empty(host_pool.shape, dtype=torch.float64, device=dev)     # filled slice by slice below
    if host_pool is None:
        scans = _to_dev(scans, torch.float64, dev)
    if scans.dim() != 3 or scans.shape[2] not in (2, 3):
        raise ValueError("scans must have shape [S, n_max, 2|3]")
    n_scans, n_max, dim = scans.shape
    if (n_max * dim) % 2:
        raise ValueError("n_max*dim must be even (16-byte aligned rows); use pack_scans")
    counts = _to_dev(counts, torch.int32, dev)
    pairs_host = None
    if host_pool is not None:
        pairs_host = (pairs.numpy() if isinstance(pairs, torch.Tensor) else np.asarray(pairs)).reshape(-1, 2)
    pairs = _to_dev(pairs, torch.int32, dev).reshape(-1, 2)
    n_pairs = pairs.shape[0]
    planar_ok = True
    if init_transforms is None:
        init = torch.eye(4, dtype=torch.float64, device=dev).repeat(n_pairs, 1, 1)
    else:
        if dim == 2 and planar_fallback and n_pairs > 0:
            # checked where the transforms already live: free for host inputs, one small device
            # reduction (and a sync) for device inputs -- pass planar_fallback=False to skip it
            # when the caller guarantees TSE3_from_TSE2-shaped inits (status 5 flags violations).
            host = init_transforms if isinstance(init_transforms, torch.Tensor) else \
                torch.as_tensor(np.asarray(init_transforms, dtype=np.float64))
            planar_ok = _keeps_plane(host.reshape(-1, 4, 4))
        init = _to_dev(init_transforms, torch.float64, dev).reshape(-1, 4, 4)
        if init.shape[0] != n_pairs:
            raise ValueError("init_transforms must have one 4x4 per pair")
    eps = _per_pair(epsilon, n_pairs, torch.float64, dev)
    norm = _per_pair(normalize_error, n_pairs, torch.uint8, dev)

    T = torch.empty((n_pairs, 4, 4), dtype=torch.float64, device=dev)
    err = torch.empty((n_pairs,), dtype=torch.float64, device=dev)
    iters = torch.empty((n_pairs,), dtype=torch.int32, device=dev)
    status = torch.empty((n_pairs,), dtype=torch.int32, device=dev)
    corr = torch.empty((n_pairs, n_max), dtype=torch.int32, device=dev) if return_correspondences else None
    if n_pairs == 0:
        return IcpBatchResult(T, err, iters, status, corr)

    nbytes = ctypes.c_size_t()
    _lib.check(lib.b200icp_workspace_bytes(n_pairs, n_max, dim, ctypes.byref(nbytes)), "b200icp_workspace_bytes")
    ws = torch.empty((nbytes.value,), dtype=torch.uint8, device=dev)
    global _last_workspace
    _last_workspace = ws

    def launch(d, scan_t, pair_t, init_t, eps_t, norm_t, T_t, err_t, it_t, st_t, corr_t, nmax, ws=ws, row0=0, gate=None):
        last = row0 + pair_t.shape[0] >= n_pairs        # chunked launches: only the last one raises the flags
        fl = peer_flag if (peer_flag and last) else (0, 0, 0)
        ex, keep = _lib.make_exchange(peer_pose_ptrs, int(peer_row0) + row0, *fl)
        with torch.cuda.device(dev):
            _lib.check(lib.b200icp_run_ex(
                _ptr(scan_t), _ptr(counts), n_scans, nmax, d, _ptr(pair_t), _ptr(init_t), _ptr(eps_t),
                _ptr(norm_t), pair_t.shape[0], int(max_iters), float(stopping_thresh), _ptr(T_t),
                _ptr(err_t), _ptr(it_t), _ptr(st_t), _ptr(corr_t), _ptr(ws), ws.numel(),
                _stream_ptr(), _ptr(gate), None if ex is None else ctypes.byref(ex), int(flags)), "b200icp_run_ex")

    if host_pool is not None and dim == 2 and n_max <= 2048 and (planar_ok or not planar_fallback):
        _copy_pool_gated(host_pool, scans, dev,
                         lambda flag: launch(dim, scans, pairs, init, eps, norm, T, err, iters, status, corr, n_max,
                                             gate=flag))
        # the gated kernel gives a pair up (status 7) rather than hang if its scans never arrive:
        # that must not pass for a result (workspace word 5, include/b200icp.h)
        torch.cuda.current_stream(dev).synchronize()
        if int(ws[20:24].cpu().numpy().view(np.uint32)[0]) != 0:
            raise RuntimeError("b200icp_run_gated: the scan pool did not arrive within the wait limit; "
                               "pairs with status 7 (not_resident) were not aligned")
        return IcpBatchResult(T, err, iters, status, corr)
    if host_pool is not None:
        _copy_pool_pipelined(host_pool, scans, pairs_host, dev,
                             None if (dim == 2 and planar_fallback and not planar_ok) else
                             (lambda lo, hi, ws_c: launch(dim, scans, pairs[lo:hi], init[lo:hi], eps[lo:hi], norm[lo:hi],
                                                          T[lo:hi], err[lo:hi], iters[lo:hi], status[lo:hi],
                                                          None if corr is None else corr[lo:hi], n_max, ws=ws_c, row0=lo)))
        if not (dim == 2 and planar_fallback and not planar_ok):
            return IcpBatchResult(T, err, iters, status, corr)
    if dim == 2 and planar_fallback and not planar_ok:
        # An initial transform that leaves the plane: the reference lifts the scans to z = 0 and
        # carries on in 3-D (icp.py:153-156), so the whole batch goes through the general kernel
        # on a z-padded copy of the pool.
        scans3 = torch.zeros((n_scans, n_max, 3), dtype=torch.float64, device=dev)
        scans3[:, :, :2] = scans
        _lib.check(lib.b200icp_workspace_bytes(n_pairs, n_max, 3, ctypes.byref(nbytes)), "b200icp_workspace_bytes")
        ws3 = torch.empty((nbytes.value,), dtype=torch.uint8, device=dev)
        launch(3, scans3, pairs, init, eps, norm, T, err, iters, status, corr, n_max, ws=ws3)
    else:
        launch(dim, scans, pairs, init, eps, norm, T, err, iters, status, corr, n_max)
    return IcpBatchResult(T, err, iters, status, corr)


def run_icp_ranges_batch(ranges, pairs, init_transforms=None, epsilon=0.01, max_iters=2000, stopping_thresh=0.0001,
                         normalize_error=False, return_correspondences=False, range_min=0.1, range_max=30.0,
                         offset_xy=(0.13323, 0.0), peer_pose_ptrs=None, peer_row0=0, flags=0, peer_flag=None):
    """run_icp_batch over a pool of lidar RANGES [S, n_beams] (float64, or float32 when every range is
    float32-representable) instead of points: get_lidar_data (modules/localization.py:252-307) is
    fused into the planar kernel's staging (b200icp_run_ranges), bit-identical to
    scans.scans_from_ranges + run_icp_batch.  Returns (IcpBatchResult, counts [S] int32: valid beams of
    every scan a pair referenced, -1 elsewhere).  Initial transforms must keep the plane."""
    from .scans import beam_tables
    lib = _lib.load()
    dev = ranges.device if isinstance(ranges, torch.Tensor) and ranges.is_cuda else _device()
    if isinstance(ranges, torch.Tensor) and ranges.dtype == torch.float32:
        r = ranges.to(dev).contiguous()
    elif isinstance(ranges, np.ndarray) and ranges.dtype == np.float32:
        r = torch.from_numpy(np.ascontiguousarray(ranges)).to(dev)
    else:
        r = _to_dev(ranges, torch.float64, dev)
    if r.dim() != 2 or r.shape[1] == 0 or r.shape[1] > 2048:
        raise ValueError("ranges must have shape [S, n_beams] with n_beams <= 2048")
    n_scans, n_beams = r.shape
    n_max = n_beams + (n_beams & 1)
    c, s_ = beam_tables(n_beams)
    tab = torch.as_tensor(np.stack([c, s_])).to(dev)
    pairs = _to_dev(pairs, torch.int32, dev).reshape(-1, 2)
    n_pairs = pairs.shape[0]
    if init_transforms is None:
        init = torch.eye(4, dtype=torch.float64, device=dev).repeat(n_pairs, 1, 1)
    else:
        init = _to_dev(init_transforms, torch.float64, dev).reshape(-1, 4, 4)
        if init.shape[0] != n_pairs:
            raise ValueError("init_transforms must have one 4x4 per pair")
    eps = _per_pair(epsilon, n_pairs, torch.float64, dev)
    norm = _per_pair(normalize_error, n_pairs, torch.uint8, dev)
    T = torch.empty((n_pairs, 4, 4), dtype=torch.float64, device=dev)
    err = torch.empty((n_pairs,), dtype=torch.float64, device=dev)
    iters = torch.empty((n_pairs,), dtype=torch.int32, device=dev)
    status = torch.empty((n_pairs,), dtype=torch.int32, device=dev)
    corr = torch.empty((n_pairs, n_max), dtype=torch.int32, device=dev) if return_correspondences else None
    counts = torch.full((n_scans,), -1, dtype=torch.int32, device=dev)
    if n_pairs == 0:
        return IcpBatchResult(T, err, iters, status, corr), counts
    nbytes = ctypes.c_size_t()
    _lib.check(lib.b200icp_workspace_bytes(n_pairs, n_max, 2, ctypes.byref(nbytes)), "b200icp_workspace_bytes")
    ws = torch.empty((nbytes.value,), dtype=torch.uint8, device=dev)
    global _last_workspace
    _last_workspace = ws
    ex, keep = _lib.make_exchange(peer_pose_ptrs, int(peer_row0), *(peer_flag or (0, 0, 0)))
    with torch.cuda.device(dev):
        _lib.check(lib.b200icp_run_ranges(
            _ptr(r), 1 if r.dtype == torch.float32 else 0, n_scans, n_beams, _ptr(tab[0]), _ptr(tab[1]),
            float(range_min), float(range_max), float(offset_xy[0]), float(offset_xy[1]), _ptr(pairs), _ptr(init),
            _ptr(eps), _ptr(norm), n_pairs, int(max_iters), float(stopping_thresh), _ptr(T), _ptr(err), _ptr(iters),
            _ptr(status), _ptr(corr), _ptr(counts), _ptr(ws), ws.numel(), _stream_ptr(), None,
            None if ex is None else ctypes.byref(ex), int(flags)), "b200icp_run_ranges")
    return IcpBatchResult(T, err, iters, status, corr), counts


_last_workspace = None


def swept_tiles() -> int:
    """(32-source row x 32-target block) tiles the planar kernel swept in the most recent
    single-launch run_icp_batch call (workspace bytes 8..15, include/b200icp.h); x 1024 = candidate
    distance evaluations including padding lanes.  Synchronises."""
    if _last_workspace is None:
        raise RuntimeError("no run_icp_batch call yet")
    return int(_last_workspace[8:16].cpu().numpy().view(np.uint64)[0])


def pack_poses(T: torch.Tensor, error: Optional[torch.Tensor] = None) -> torch.Tensor:
    """[B,4,4] transforms (+ errors) -> [B,4] float64 rows (x, y, theta, error): pose_from_T of
    modules/utils.py:196-210; the payload of the multi-GPU all-gather."""
    lib = _lib.load()
    T = T.contiguous()
    out = torch.empty((T.shape[0], 4), dtype=torch.float64, device=T.device)
    with torch.cuda.device(T.device):
        _lib.check(lib.b200icp_pack_poses(_ptr(T), _ptr(error), T.shape[0], _ptr(out), _stream_ptr()),
                   "b200icp_pack_poses")
    return out


def shard_range(n_items: int, rank: int, world: int):
    """Contiguous block partition used for pairs: rank r owns [lo, hi); every rank gets
    ceil(n/world) slots so that the all-gather is uniform (tail slots are padding)."""
    per = -(-n_items // world)
    lo = min(rank * per, n_items)
    return lo, min(lo + per, n_items), per


_PEER_BUFFERS = {}
_FLAG_WORDS = 16          # one uint32 arrival flag per rank (<= 16 ranks), per row slot


class PeerGather:
    """Symmetric-memory gather buffer of the fused pose exchange, shared by all ranks of a group:
    DEPTH row buffers [rows, 4] float64 used in turn (step s uses slot s % DEPTH) and one uint32 arrival
    flag per (slot, source rank).  Every rank maps every other rank's copy, so a kernel stores rows --
    and, when its last pair retires, its flag -- straight into all of them over NVLink.  Protocol of
    step s (value s + 1), with NO host-side cross-rank barrier:
      * the ICP launch stores this rank's rows into slot s % DEPTH of every rank, then flag[slot][rank] = s + 1;
      * `wait(token)` (a stream-ordered kernel) returns once flag[slot][r] >= s + 1 for every rank r: all
        rows of step s are in this rank's slot;
      * the slot is written again in step s + DEPTH.  A caller that waits right after each launch
        (run_icp_sharded: L(s) W(s) C(s) L(s+1) ...) needs DEPTH >= 2: a peer launches step s + 2 after its
        wait of step s + 1, i.e. after this rank's kernel s + 1 finished, which in stream order follows
        this rank's consumers C(s).  A caller that defers the wait by one step to hide the slowest rank
        (bench.py: L(s+1) W(s) C(s) L(s+2) ...) needs DEPTH >= 4: a peer launches step s + 4 after its wait
        of step s + 2, i.e. after this rank's kernel s + 2, which is enqueued behind C(s)."""

    DEPTH = 4

    def __init__(self, rows: int, dev, group):
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm_mem
        self.rows, self.dev = rows, dev
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        n_f64 = self.DEPTH * rows * 4 + self.DEPTH * _FLAG_WORDS // 2      # row slots + DEPTH x 16 uint32 flags
        self.buf = symm_mem.empty((n_f64,), dtype=torch.float64, device=dev)
        self.buf.zero_()
        self.hdl = symm_mem.rendezvous(self.buf, group=group if group is not None else dist.group.WORLD)
        torch.cuda.synchronize(dev)
        self.hdl.barrier(channel=0)                           # every rank's flags are zero before anyone stores
        self.step = 0
        self.timed_out = torch.zeros((1,), dtype=torch.int32, device=dev)
        self.base_ptrs = [int(q) for q in self.hdl.buffer_ptrs]

    def begin(self):
        """(peer pointers of this step's row slot, peer_flag triple, token for wait / rows_view)"""
        s = self.step
        self.step += 1
        slot = s % self.DEPTH
        ptrs = [b + slot * self.rows * 32 for b in self.base_ptrs]
        # flag offset is relative to the row pointer the kernel gets (the slot's base)
        flag_off = (self.DEPTH - slot) * self.rows * 32 + slot * 4 * _FLAG_WORDS
        return ptrs, (flag_off, self.rank, s + 1), (slot, s)

    def rows_view(self, token, rank=None):
        """[rows, 4] view of this rank's (or a peer's) row slot of the step `token` names."""
        slot = token[0] if isinstance(token, tuple) else int(token)
        if rank is None or rank == self.rank:
            return self.buf[slot * self.rows * 4:(slot + 1) * self.rows * 4].view(self.rows, 4)
        whole = self.hdl.get_buffer(rank, (self.buf.numel(),), torch.float64)
        return whole[slot * self.rows * 4:(slot + 1) * self.rows * 4].view(self.rows, 4)

    def wait(self, token):
        """Stream-ordered: every rank's rows of the step `token` names have landed in this rank's slot."""
        slot, s = token
        lib = _lib.load()
        flags_ptr = self.buf.data_ptr() + self.DEPTH * self.rows * 32 + slot * 4 * _FLAG_WORDS
        with torch.cuda.device(self.dev):
            _lib.check(lib.b200icp_exchange_wait(ctypes.c_void_p(flags_ptr), self.world, s + 1, 20000,
                                                 _ptr(self.timed_out), _stream_ptr()), "b200icp_exchange_wait")
        return self.rows_view(slot)

    def check(self):
        if int(self.timed_out.item()) != 0:
            raise RuntimeError("fused pose exchange: a rank's arrival flag did not show up within 20 s")


def _peer_gather_buffer(rows: int, dev, group):
    """Cached PeerGather for (rows, device, group)."""
    key = (rows, str(dev), id(group))
    if key not in _PEER_BUFFERS:
        _PEER_BUFFERS[key] = PeerGather(rows, dev, group)
    return _PEER_BUFFERS[key]


def run_icp_sharded(scans, counts, pairs, init_transforms=None, group=None, exchange="auto", **kw):
    """Multi-GPU entry: every rank holds the scan pool (or at least the scans its pairs use) and
    aligns its contiguous shard of `pairs`; the [B,4] (x, y, theta, error) rows of ALL pairs, in
    pair order, end up on every rank -- what the pose-graph builder needs (main.py:148-172).
    Pairs are independent, so this is the only exchange.

    exchange="peer": fused -- the ICP kernel's epilogue stores each finished row directly into
    every rank's symmetric-memory gather buffer (peer stores over NVLink / NVSwitch) and, when the
    rank's last pair retires, an arrival flag; a stream-ordered wait on the flags of all ranks
    follows.  No gather kernel, no collective, no host-side barrier (PeerGather).
    exchange="nccl": pack_poses + ONE all_gather_into_tensor (NCCL on GPUs, gloo in the CPU tests of
    the host logic).  "auto" = "peer" on CUDA with NCCL when symmetric memory is available, else "nccl".
    Returns (rows [B,4], local IcpBatchResult); with "peer" the rows alias the cached buffer and
    are valid until PeerGather.DEPTH - 1 further sharded calls of the same shape (the row slots are used in turn).
    """
    import torch.distributed as dist
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    pairs_t = pairs if isinstance(pairs, torch.Tensor) else torch.as_tensor(np.asarray(pairs))
    pairs_t = pairs_t.reshape(-1, 2)
    n = pairs_t.shape[0]
    lo, hi, per = shard_range(n, rank, world)
    init_loc = None if init_transforms is None else init_transforms[lo:hi]
    for key in ("epsilon", "normalize_error"):
        v = kw.get(key)
        if v is not None and not isinstance(v, (int, float, bool)):
            kw[key] = v[lo:hi]
    pg = None
    if world > 1 and exchange in ("auto", "peer") and torch.cuda.is_available() and dist.get_backend(group) == "nccl":
        try:
            pg = _peer_gather_buffer(world * per, _device(), group)
        except Exception:                      # symmetric memory unavailable on this build / topology
            if exchange == "peer":
                raise
    if pg is not None:
        ptrs, flag, parity = pg.begin()
        if hi - lo < per:                      # this rank's padding rows, to every peer, ahead of the launch
            for r in range(world):
                pg.rows_view(parity, r)[rank * per + (hi - lo):(rank + 1) * per] = float("nan")
        if hi > lo:
            res = run_icp_batch(scans, counts, pairs_t[lo:hi], init_loc, peer_pose_ptrs=ptrs, peer_row0=rank * per,
                                peer_flag=flag, **kw)
        else:                                  # an empty shard still announces itself
            res = run_icp_batch(scans, counts, pairs_t[lo:hi], init_loc, **kw)
            ex, keep = _lib.make_exchange(ptrs, rank * per, *flag)
            with torch.cuda.device(pg.dev):
                _lib.check(_lib.load().b200icp_exchange_signal(ctypes.byref(ex), _stream_ptr()), "b200icp_exchange_signal")
        rows = pg.wait(parity)
        return rows[:n], res
    res = run_icp_batch(scans, counts, pairs_t[lo:hi], init_loc, **kw)
    local = torch.full((per, 4), float("nan"), dtype=torch.float64, device=res.T.device)
    if hi > lo:
        local[:hi - lo] = pack_poses(res.T, res.error)
    if world == 1:
        return local[:n], res
    gathered = torch.empty((world * per, 4), dtype=torch.float64, device=local.device)
    dist.all_gather_into_tensor(gathered, local, group=group)
    return gathered[:n], res


def nearest_batch(scans, counts, pairs, return_dist2=False):
    """Exact nearest neighbour of every point of scan pairs[b,0] in scan pairs[b,1]
    (get_correspondences / KDTree.query, modules/icp.py:29-41), no transform."""
    lib = _lib.load()
    dev = scans.device if isinstance(scans, torch.Tensor) and scans.is_cuda else _device()
    scans = _to_dev(scans, torch.float64, dev)
    n_scans, n_max, dim = scans.shape
    counts = _to_dev(counts, torch.int32, dev)
    pairs = _to_dev(pairs, torch.int32, dev).reshape(-1, 2)
    n_pairs = pairs.shape[0]
    corr = torch.empty((n_pairs, n_max), dtype=torch.int32, device=dev)
    d2 = torch.empty((n_pairs, n_max), dtype=torch.float64, device=dev) if return_dist2 else None
    nbytes = ctypes.c_size_t()
    _lib.check(lib.b200icp_workspace_bytes(n_pairs, n_max, dim, ctypes.byref(nbytes)), "b200icp_workspace_bytes")
    ws = torch.empty((nbytes.value,), dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        _lib.check(lib.b200icp_nearest(_ptr(scans), _ptr(counts), n_scans, n_max, dim, _ptr(pairs), n_pairs,
                                       _ptr(corr), _ptr(d2), _ptr(ws), ws.numel(), _stream_ptr()),
                   "b200icp_nearest")
    return (corr, d2) if return_dist2 else corr


def voxel_downsample_device(points: torch.Tensor, voxel_size: float) -> torch.Tensor:
    """Device voxel_downsample (modules/icp.py:4-27) on an [N,3] float64 CUDA tensor."""
    lib = _lib.load()
    dev = points.device
    points = points.to(torch.float64).contiguous()
    n = points.shape[0]
    nbytes = ctypes.c_size_t()
    with torch.cuda.device(dev):
        _lib.check(lib.b200icp_voxel_workspace_bytes(n, ctypes.byref(nbytes)), "b200icp_voxel_workspace_bytes")
        ws = torch.empty((nbytes.value,), dtype=torch.uint8, device=dev)
        out = torch.empty((n, 3), dtype=torch.float64, device=dev)
        n_out = torch.zeros((1,), dtype=torch.int32, device=dev)
        _lib.check(lib.b200icp_voxel_downsample(_ptr(points), n, float(voxel_size), _ptr(out), _ptr(n_out),
                                                _ptr(ws), ws.numel(), _stream_ptr()), "b200icp_voxel_downsample")
    m = int(n_out.item())
    if m < 0:
        raise ValueError("voxel grid exceeds 2^21 cells along an axis; use a larger voxel_size")
    return out[:m]


def ffma_peak_tflops(seconds: float = 0.3) -> dict:
    """Measured FP32 FFMA issue rate of the current device (the NN kernel's roofline
    denominator): lane-FFMA/s and the equivalent TFLOP/s (2 flops per FFMA)."""
    lib = _lib.load()
    dev = _device()
    props = torch.cuda.get_device_properties(dev)
    blocks = props.multi_processor_count * 8
    sink = torch.empty((blocks * 256,), dtype=torch.float32, device=dev)
    iters = 4096
    start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    best = 0.0
    for rep in range(6):
        start.record()
        _lib.check(lib.b200icp_ffma_probe(_ptr(sink), blocks, iters, _stream_ptr()), "b200icp_ffma_probe")
        stop.record()
        stop.synchronize()
        ms = start.elapsed_time(stop)
        rate = blocks * 256 * iters * 64 / (ms * 1e-3)
        if rep > 0:
            best = max(best, rate)
        if ms < seconds * 1e3 / 8 and rep == 0:
            iters *= max(1, int(seconds * 1e3 / 8 / max(ms, 1e-3)))
    return {"lane_ffma_per_s": best, "tflops": 2 * best / 1e12, "sms": props.multi_processor_count}
