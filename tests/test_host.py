"""CPU-only tests of the host side: the C-ABI library loads and exports every symbol the
header declares, the synthetic generators are deterministic, the fp32 slack bound used by the
kernels holds, and the multi-rank sharding logic (gloo, world_size 2)."""
import ctypes
import os
import re
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    from lidar_slam_b200 import _lib, build
    build.build()
    header = open(os.path.join(ROOT, "include", "b200icp.h")).read()
    declared = set(re.findall(r"\b(b200icp_[a-z0-9_]+)\s*\(", header))
    assert declared == set(_lib.SIGNATURES), declared ^ set(_lib.SIGNATURES)
    lib = _lib.load()
    for name in declared:
        assert hasattr(lib, name), name
    assert lib.b200icp_version() == 100
    assert b"workspace" in lib.b200icp_error_string(-3)
    n = ctypes.c_size_t()
    assert lib.b200icp_workspace_bytes(10, 1081, 2, ctypes.byref(n)) == 0 and n.value >= 4
    assert lib.b200icp_workspace_bytes(10, 1081, 5, ctypes.byref(n)) == -1     # bad dim


def test_product_path_fails_loudly_without_gpu():
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import lidar_slam_b200 as eng
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        eng.run_icp(np.zeros((5, 2)), np.ones((5, 2)))


def test_host_entries_validate_before_touching_the_device_and_fail_loudly_without_one():
    """hostapi: page-locked buffers fall back to pageable ones when cudaHostAlloc is unavailable, the
    submit wrapper rejects arrays the library could not read asynchronously, and without a GPU every
    entry raises (no CPU fallback); run_static reports its size limit without a device, too."""
    from lidar_slam_b200 import _lib, hostapi
    a = hostapi.pinned_empty((7, 3), np.float32)
    assert a.shape == (7, 3) and a.dtype == np.float32 and a.flags.c_contiguous
    a[:] = 1.5
    ws = hostapi.HostRangesWorkspace(4, 5, 1082)
    assert ws.T.shape == (4, 4, 4) and ws.counts.shape == (5,)
    c, s_ = ws.beam_tables(1081)
    c0, s0 = hostapi.beam_tables(1081)
    assert np.array_equal(c, c0) and np.array_equal(s_, s0)
    eps, norm = ws.per_pair(1e-3, True)
    assert eps.shape == (4,) and float(eps[2]) == 1e-3 and norm.dtype == np.uint8 and int(norm[3]) == 1
    ranges = np.full((5, 1081), 4.0)
    pairs = np.array([[1, 0], [2, 1], [3, 2], [4, 3]], np.int32)
    init = np.broadcast_to(np.eye(4), (4, 4, 4)).copy()
    with pytest.raises(ValueError, match="pairs"):
        hostapi.submit_icp_ranges_host(ranges, pairs.astype(np.int64), init, ws)
    with pytest.raises(ValueError, match="ranges"):
        hostapi.submit_icp_ranges_host(ranges[:, ::2], pairs, init, ws)
    with pytest.raises(ValueError, match="another batch shape"):
        hostapi.submit_icp_ranges_host(ranges, pairs[:3].copy(), init[:3].copy(), ws)
    lib = _lib.load()
    assert lib.b200icp_wait_host(-1) == -1 and lib.b200icp_wait_host(5 << 8) == -1
    assert b"in flight" in lib.b200icp_error_string(-7)
    if not torch.cuda.is_available():
        with pytest.raises(RuntimeError, match="no CPU fallback"):
            hostapi.submit_icp_ranges_host(ranges, pairs, init, ws)
        with pytest.raises(RuntimeError, match="no CPU fallback"):
            hostapi.run_icp_ranges_host(ranges, pairs, init)
        # the size limit of the allocation-free entry is checked before the device is
        vp = ctypes.c_void_p
        assert lib.b200icp_run_static(vp(8), vp(8), 2, 20000, 3, vp(8), vp(8), vp(8), vp(8), 512, 10, 1e-4,
                                      vp(8), None, None, None, None, None) == _lib.ERR_STATIC_SCRATCH


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "lidar_slam_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in text and "from oracle" not in text, f


def test_generators_are_deterministic_and_shaped_like_hokuyo():
    from lidar_slam_b200 import synth
    a = synth.odometry_batch(5, seed=3)
    b = synth.odometry_batch(5, seed=3)
    assert torch.equal(a["scans"], b["scans"]) and torch.equal(a["init"], b["init"])
    assert a["scans"].shape == (6, 1081, 2) and a["scans"].dtype == torch.float64
    c = a["counts"].numpy()
    assert np.all(c > 1040) and np.all(c <= 1081)
    r = np.hypot(a["scans"][0, :c[0], 0].numpy() - synth.LIDAR_FORWARD, a["scans"][0, :c[0], 1].numpy())
    assert r.min() >= 0.1 and r.max() <= 30.0
    assert np.allclose(r * 1000, np.round(r * 1000), atol=1e-3)        # mm-quantised ranges
    assert np.array_equal(a["pairs"].numpy()[:, 0], np.arange(1, 6))
    init = a["init"].numpy()
    assert np.all(init[:, 2, 2] == 1) and np.all(init[:, 3, 3] == 1) and np.all(init[:, 2, :2] == 0)


def _slack(d_hi, a_bound):
    """float32 restatement of slack() in csrc/icp_common.cuh."""
    u = np.float32(2.0 ** -24)
    s = np.sqrt(np.float32(d_hi))
    a_s = np.float32(a_bound) + s
    return np.float32(8) * u * a_s * s + np.float32(8) * u * np.float32(d_hi) + np.float32(8) * u * u * a_s * a_s + np.float32(1.1754944e-38)


@pytest.mark.parametrize("dim,scale", [(2, 5.0), (2, 30.0), (3, 1.0), (3, 1000.0), (2, 1e-3)])
def test_fp32_slack_bounds_the_fp64_distance(dim, scale):
    """|d32 - d64| <= slack(d32, A): the inequality the exact-NN scheme rests on."""
    rng = np.random.default_rng(dim * 7 + int(scale))
    for spread in (1e-4, 1e-2, 1.0):
        t = rng.uniform(-scale, scale, size=(4000, dim))
        q = t + rng.normal(scale=spread * scale, size=t.shape)
        q32, t32 = q.astype(np.float32), t.astype(np.float32)
        diff = q32 - t32
        d32 = np.zeros(len(t), np.float32)
        for a in range(dim):
            d32 = d32 + diff[:, a] * diff[:, a]
        d64 = ((q - t) ** 2).sum(1)
        a_bound = np.abs(q32).max(1) + np.abs(t32).max()
        bound = _slack(d32, a_bound).astype(np.float64)
        assert np.all(np.abs(d32.astype(np.float64) - d64) <= bound)
        assert np.all(np.abs(d32.astype(np.float64) - d64) <= 0.5 * bound)    # the documented factor 2


def test_box_lower_bound_never_exceeds_the_fp32_candidate_distance():
    """The pruned sweep (icp2d.cuh, box_lb) skips a target block when lb(row box, block box) exceeds
    the candidate threshold.  lb is computed with the candidate distance's own fp32 operations on the
    per-axis gaps; rounding to nearest is monotone, so lb <= d32(q, t) for EVERY q in the row box and
    t in the block box, with no error term.  Emulated here in numpy float32 (FMA in float64 rounded
    once), on boxes that overlap, touch and lie far apart, at small and large coordinates."""
    f32 = np.float32

    def fma32(a, b, c):        # one rounding, like FFMA (exact product and sum fit float64 for fp32 inputs)
        return (a.astype(np.float64) * b.astype(np.float64) + c.astype(np.float64)).astype(f32)

    def d32(qx, qy, tx, ty):
        dx, dy = (qx - tx).astype(f32), (qy - ty).astype(f32)
        return fma32(dx, dx, (dy * dy).astype(f32))

    rng = np.random.default_rng(5)
    for scale, spread in ((1.0, 0.3), (30.0, 2.0), (30.0, 1e-3), (4000.0, 50.0), (1e-3, 1e-4)):
        for _ in range(200):
            q = (rng.uniform(-scale, scale, 2) + rng.normal(scale=spread, size=(32, 2))).astype(f32)
            t = (rng.uniform(-scale, scale, 2) + rng.normal(scale=spread, size=(32, 2))).astype(f32)
            if rng.random() < 0.3:
                t = (q[rng.integers(0, 32, 32)] + rng.normal(scale=spread * 1e-3, size=(32, 2))).astype(f32)
            qb = (q[:, 0].min(), q[:, 0].max(), q[:, 1].min(), q[:, 1].max())
            tb = (t[:, 0].min(), t[:, 0].max(), t[:, 1].min(), t[:, 1].max())
            ex = np.maximum(np.maximum(f32(tb[0]) - f32(qb[1]), f32(qb[0]) - f32(tb[1])), f32(0)).astype(f32)
            ey = np.maximum(np.maximum(f32(tb[2]) - f32(qb[3]), f32(qb[2]) - f32(tb[3])), f32(0)).astype(f32)
            lb = fma32(np.array(ex), np.array(ex), np.array((ey * ey), dtype=f32))
            d = d32(q[:, None, 0], q[:, None, 1], t[None, :, 0], t[None, :, 1])
            assert lb <= d.min(), (scale, spread, float(lb), float(d.min()))


def _candidate_threshold(trunc_min, a_bound):
    """float32 restatement of candidate_threshold() in csrc/icp_common.cuh (trunc_rel = 2^-16)."""
    f32 = np.float32
    hi = (trunc_min.astype(np.float64) * 2.0 ** -16 + trunc_min.astype(np.float64)).astype(f32)
    thr = (hi + f32(4) * _slack(hi, a_bound)).astype(f32)
    return (thr.astype(np.float64) * (4 * 2.0 ** -24) + thr.astype(np.float64)).astype(f32)


def _trunc(d):
    return (d.view(np.uint32) & np.uint32(0xffffffc0)).view(np.float32)


@pytest.mark.parametrize("kind", ["lidar", "lidar_shifted", "clusters", "shuffled", "two_blocks"])
def test_pruned_search_emulation_equals_float64_exhaustive_search(kind):
    """numpy restatement of the pruned NN search of csrc/icp2d.cuh (sweep_row_pruned + resolution), in
    float32 where the kernel is float32: boxes, lower bounds, two seed blocks, the threshold of the
    current best, the blocks that survive it, then candidates <= thr of the best block re-evaluated in
    float64 (all targets if a second swept block is within thr).  The result must be the float64
    exhaustive argmin (lowest index on ties) for every query -- the property the GPU tests check
    against the device; this test checks the ALGORITHM on the CPU, and that blocks are really skipped."""
    f32 = np.float32
    rng = np.random.default_rng({"lidar": 1, "lidar_shifted": 2, "clusters": 3, "shuffled": 4, "two_blocks": 5}[kind])
    if kind in ("lidar", "lidar_shifted", "shuffled"):
        from lidar_slam_b200 import synth
        b = synth.odometry_batch(2, seed=31)
        c = b["counts"].numpy()
        tgt, src = b["scans"][0, :c[0]].numpy().copy(), b["scans"][1, :c[1]].numpy().copy()
        if kind == "lidar_shifted":
            tgt += np.array([310.5, -127.25]); src += np.array([310.5, -127.25])
        if kind == "shuffled":
            tgt = tgt[rng.permutation(len(tgt))]
    elif kind == "clusters":
        centres = rng.uniform(-20, 20, size=(9, 2))
        tgt = np.concatenate([cc + rng.normal(scale=0.05, size=(40, 2)) for cc in centres])
        tgt[31] = tgt[32]
        src = tgt[rng.permutation(len(tgt))[:200]] + rng.normal(scale=0.003, size=(200, 2))
        src[0] = tgt[31]
    else:
        tgt = rng.uniform(-1, 1, size=(50, 2))
        src = rng.uniform(-1, 1, size=(70, 2))
    nt, ns = len(tgt), len(src)
    nblk = (nt + 31) // 32
    t32, q32 = tgt.astype(f32), src.astype(f32)
    tmax = f32(np.abs(t32).max()) * (f32(1) + f32(4 * 2.0 ** -24))

    def fma32(a, b_, c_):
        return (a.astype(np.float64) * b_.astype(np.float64) + c_.astype(np.float64)).astype(f32)

    def d32(q, t):          # [nq, ntargets]
        dx = (q[:, None, 0] - t[None, :, 0]).astype(f32)
        dy = (q[:, None, 1] - t[None, :, 1]).astype(f32)
        return fma32(dx, dx, (dy * dy).astype(f32))

    boxes = np.array([[t32[k * 32:(k + 1) * 32, 0].min(), t32[k * 32:(k + 1) * 32, 0].max(),
                       t32[k * 32:(k + 1) * 32, 1].min(), t32[k * 32:(k + 1) * 32, 1].max()] for k in range(nblk)], f32)
    d64_all = ((src[:, None, :] - tgt[None, :, :]) ** 2).sum(-1)
    want = d64_all.argmin(1)                       # numpy argmin: lowest index on ties
    swept_total = 0
    for r0 in range(0, ns, 32):
        q = q32[r0:r0 + 32]
        qb = (q[:, 0].min(), q[:, 0].max(), q[:, 1].min(), q[:, 1].max())
        ex = np.maximum(np.maximum(boxes[:, 0] - qb[1], qb[0] - boxes[:, 1]), f32(0)).astype(f32)
        ey = np.maximum(np.maximum(boxes[:, 2] - qb[3], qb[2] - boxes[:, 3]), f32(0)).astype(f32)
        lb = fma32(ex, ex, (ey * ey).astype(f32))
        key = (lb.view(np.uint32) & np.uint32(0xffffffc0)) | np.arange(nblk, dtype=np.uint32)
        seeds = list(np.argsort(key)[:2])
        blockmin = {}

        def sweep(k):
            blockmin[k] = d32(q, t32[k * 32:(k + 1) * 32]).min(1)
        for k in seeds:
            sweep(k)

        def best_second():
            ks = sorted(blockmin)
            pk = np.stack([_trunc(blockmin[k]).view(np.uint32) | np.uint32(k) for k in ks], 1)
            order = np.sort(pk, 1)
            return order[:, 0], (order[:, 1] if len(ks) > 1 else np.full(len(q), 0xffffffff, np.uint32))
        best, _ = best_second()
        a_bound = np.maximum(np.abs(q[:, 0]), np.abs(q[:, 1])) * (f32(1) + f32(4 * 2.0 ** -24)) + tmax
        thr = _candidate_threshold((best & np.uint32(0xffffffc0)).view(f32), a_bound)
        bound = thr.view(np.uint32).max()
        for k in range(nblk):
            if k not in blockmin and not (lb[k:k + 1].view(np.uint32)[0] > bound):
                sweep(k)
        swept_total += len(blockmin)
        best, second = best_second()
        thr = _candidate_threshold((best & np.uint32(0xffffffc0)).view(f32), a_bound)
        for i in range(len(q)):
            blk = int(best[i] & np.uint32(63))
            sval = (second[i:i + 1] & np.uint32(0xffffffc0)).view(f32)[0]
            if not (sval > thr[i]):
                cand = np.arange(nt)[~(d32(q[i:i + 1], t32)[0] > thr[i])]          # fallback: all targets
            else:
                idx = np.arange(blk * 32, min(nt, blk * 32 + 32))
                dd = d32(q[i:i + 1], t32[idx])[0]
                groups = {g for g in range(len(idx) // 4 + 1) if len(dd[4 * g:4 * g + 4]) and dd[4 * g:4 * g + 4].min() <= thr[i]}
                cand = np.array([j for j in idx if (j - blk * 32) // 4 in groups])
            got = cand[np.argmin(d64_all[r0 + i, cand])]
            assert got == want[r0 + i], (kind, r0 + i, got, want[r0 + i])
    rows = (ns + 31) // 32
    if kind in ("lidar", "lidar_shifted"):
        assert swept_total < 0.2 * rows * nblk            # ordered scans: most blocks are skipped


def test_host_kabsch3_matches_lapack(tmp_path):
    """kabsch3.cuh (compiled for the host with g++) against the reference's SVD recipe."""
    src = tmp_path / "hk.cpp"
    src.write_text('#include "%s"\nextern "C" void hk(const double* C, double* R)'
                   '{ b200icp::kabsch_rotation_3x3(C, R); }\n'
                   % os.path.join(ROOT, "lidar_slam_b200", "csrc", "kabsch3.cuh"))
    so = tmp_path / "hk.so"
    import subprocess
    subprocess.run(["g++", "-O2", "-shared", "-fPIC", "-o", str(so), str(src)], check=True)
    lib = ctypes.CDLL(str(so))
    from oracle import icp_oracle as orc
    rng = np.random.default_rng(1)
    worst = 0.0
    for k in range(3000):
        n = int(rng.integers(4, 40))
        X = rng.normal(size=(n, 3)) * rng.uniform(0.05, 3, 3)
        A = np.linalg.qr(rng.normal(size=(3, 3)))[0]
        Y = X @ A.T + rng.normal(scale=0.02, size=X.shape)
        if k % 4 == 0:
            X[:, 2] = 0
            Y[:, 2] = 0
        C = np.ascontiguousarray((X - X.mean(0)).T @ (Y - Y.mean(0)))
        s = np.linalg.svd(C, compute_uv=False)
        if s[1] < 1e-6 * s[0] or s[1] - s[2] < 1e-3 * s[0]:
            continue
        R = np.zeros((3, 3))
        lib.hk(C.ctypes.data_as(ctypes.c_void_p), R.ctypes.data_as(ctypes.c_void_p))
        worst = max(worst, np.abs(R - orc.rigid_fit(X, Y)[:3, :3]).max())
    assert worst < 1e-11, worst


def test_shard_range_partitions_every_pair_once():
    from lidar_slam_b200.engine import shard_range
    for n in (0, 1, 7, 4955, 5450):
        for world in (1, 2, 3, 8):
            seen = []
            for r in range(world):
                lo, hi, per = shard_range(n, r, world)
                assert hi - lo <= per
                seen += list(range(lo, hi))
            assert seen == list(range(n))


def _sharded_worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch.distributed as dist
    from lidar_slam_b200 import engine
    from _cpu_standin import install_cpu_standin
    install_cpu_standin(engine)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    scans, counts, pairs, init = _toy_problem()
    poses, _ = engine.run_icp_sharded(scans, counts, pairs, init, epsilon=1e-3)
    np.save(os.path.join(out_dir, f"poses{rank}.npy"), poses.numpy())
    dist.destroy_process_group()


def _toy_problem():
    from lidar_slam_b200 import synth
    b = synth.odometry_batch(5, seed=8)
    scans, counts = synth.thin_scans(b["scans"], b["counts"], seed=1, lo=60, hi=90)
    return scans, counts, b["pairs"], b["init"]


def test_sharded_all_gather_world2_gloo(tmp_path):
    """run_icp_sharded with 2 gloo ranks == the single-rank result, in pair order, on every
    rank.  The device aligner is replaced by a CPU stand-in (tests/_cpu_standin.py) because
    only the sharding / gather logic is under test here."""
    import torch.multiprocessing as mp
    port = 29500 + os.getpid() % 2000
    mp.spawn(_sharded_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from lidar_slam_b200 import engine
    from _cpu_standin import install_cpu_standin, cpu_align
    scans, counts, pairs, init = _toy_problem()
    want = cpu_align(scans, counts, pairs, init, epsilon=1e-3)
    p0 = np.load(tmp_path / "poses0.npy")
    p1 = np.load(tmp_path / "poses1.npy")
    assert p0.shape == (5, 4) and np.array_equal(p0, p1)
    np.testing.assert_allclose(p0, want, rtol=0, atol=1e-12)


def test_c_abi_argument_validation_without_a_device():
    """Every entry rejects bad arguments with B200ICP_ERR_BAD_ARG / _WORKSPACE / _TOO_LARGE before it
    touches a device (no compute call is made here), and the new product modules fail loudly on a
    box without a GPU."""
    from lidar_slam_b200 import _lib
    lib = _lib.load()
    n = ctypes.c_size_t()
    null = ctypes.c_void_p(0)
    one = ctypes.c_void_p(64)            # never dereferenced: the calls fail on other arguments first
    assert lib.b200icp_run(null, one, 1, 2, 2, one, one, one, one, 1, 5, 1e-4, one, one, one, one, null, one, 256, null) == -1
    assert lib.b200icp_run(one, one, 1, 2, 2, one, one, one, one, 1, -1, 1e-4, one, one, one, one, null, one, 256, null) == -1
    assert lib.b200icp_run_exchange(one, one, 1, 2, 2, one, one, one, one, 1, 5, 1e-4, one, one, one, one, null, one, 256,
                                    null, null, 3, 0) == -1                      # peers announced but no array
    # gated entry: the counter of resident scans is mandatory, and only the planar kernel waits on one
    assert lib.b200icp_run_gated(one, one, 1, 2, 2, one, one, one, one, 1, 5, 1e-4, one, one, one, one, null, one, 256,
                                 null, null, null, 0, 0) == -1                   # no ready_scans
    assert lib.b200icp_run_gated(one, one, 1, 2, 3, one, one, one, one, 1, 5, 1e-4, one, one, one, one, null, one, 256,
                                 null, one, null, 0, 0) == -1                    # 3-D: general kernel
    assert lib.b200icp_run_gated(one, one, 1, 4096, 2, one, one, one, one, 1, 5, 1e-4, one, one, one, one, null, one,
                                 256, null, one, null, 0, 0) == -1               # planar but n_max > 2048
    assert lib.b200icp_run_gated(one, one, 1, 2, 2, one, one, one, one, 1, 5, 1e-4, one, one, one, one, null, one, 256,
                                 null, one, null, 2, 0) == -1                    # peers announced but no array
    # host-buffer entry: bad arguments, the empty pair list (nothing to do), and no device here
    assert lib.b200icp_run_host(null, one, 1, 2, 2, one, one, one, one, 1, 5, 1e-4, one, one, one, one, null) == -1
    assert lib.b200icp_run_host(one, one, 1, 3, 3, one, one, one, one, 1, 5, 1e-4, one, one, one, one, null) == -1  # odd row
    assert lib.b200icp_run_host(one, one, 1, 2, 2, one, one, one, one, 0, 5, 1e-4, one, one, one, one, null) == 0
    import torch
    if not torch.cuda.is_available():
        z = np.zeros((1, 2, 2)); c = np.ones(1, dtype=np.int32); pr = np.zeros((1, 2), dtype=np.int32)
        from lidar_slam_b200 import hostapi
        with pytest.raises(RuntimeError):
            hostapi.run_icp_batch_host(z, c, pr)
    assert lib.b200icp_run_exchange(one, one, 1, 2, 2, one, one, one, one, 1, 5, 1e-4, one, one, one, one, null, one, 256,
                                    null, one, 17, 0) == -1                      # more than 16 peers
    assert lib.b200icp_scans_from_ranges(one, 1, 1081, one, one, 0.1, 30.0, 0.1, 0.0, one, one, 1081, null) == -1   # odd n_max
    assert lib.b200icp_range_filter_workspace_bytes(ctypes.byref(n)) == 0 and n.value > 0
    assert lib.b200icp_range_filter(one, one, 1, 1082, 3.0, one, null, 0, null) == -3
    assert lib.b200icp_ogm_workspace_bytes(100, 100, 1082, 1, ctypes.byref(n)) == 0 and n.value > 100 * 100 * 4
    assert lib.b200icp_ogm_build(one, 100, 100, 0.0, 0.0, 0.0, one, one, one, one, 1, 1082, 1, one, 1 << 30, null) == -1
    assert lib.b200icp_ogm_build(one, 100, 100, 0.05, 0.0, 0.0, one, one, one, one, 1, 40000, 1, one, 1 << 30, null) == -1
    assert lib.b200icp_ogm_build(one, 100, 100, 0.05, 0.0, 0.0, one, one, one, one, 1, 1082, 1, one, 16, null) == -3
    assert lib.b200icp_texture_paint(one, 10, 10, 0.05, 0.0, 0.0, one, one, 1, 2048, 2048, 0, one, one, one, one, null) == -2
    assert lib.b200icp_texture_paint(one, 10, 10, 0.05, 0.0, 0.0, one, one, 1, 48, 64, -1, one, one, one, one, null) == -1
    assert lib.b200icp_texture_resolve(null, null, one, 10, 10, null) == -1
    if not torch.cuda.is_available():
        from lidar_slam_b200 import ogm, scans, texture_mapping
        with pytest.raises(RuntimeError, match="no CPU fallback"):
            scans.scans_from_ranges(np.ones((2, 8)), 0.1, 30.0)
        with pytest.raises(RuntimeError, match="no CPU fallback"):
            ogm.OccupancyGridMap(0.1, 1, 1, -1, -1)
        with pytest.raises(RuntimeError, match="no CPU fallback"):
            texture_mapping.TexturePainter(10, 10, 0.1, 0, 0, np.eye(4), np.eye(3))


def test_host_binding_does_not_import_torch():
    """The numpy binding of the host-buffer entry is usable without torch (what a maintainer of the
    numpy-only reference would bind)."""
    import subprocess
    import sys
    code = ("import sys; import lidar_slam_b200.hostapi as h; import lidar_slam_b200 as p; "
            "p.pack_scans_numpy; assert 'torch' not in sys.modules, 'torch was imported'; print('ok')")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, "-c", code], cwd=root, capture_output=True, text=True, timeout=120)
    assert out.returncode == 0 and out.stdout.strip() == "ok", out.stderr


def _build_c_caller(tmp_path):
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    from lidar_slam_b200 import build
    build.build()
    exe = str(tmp_path / "abi_smoke")
    libdir = os.path.join(root, "lidar_slam_b200")
    cmd = ["gcc", "-std=c99", "-O1", "-Wall", "-Werror", "-I", os.path.join(root, "include"),
           os.path.join(root, "tests", "c", "abi_smoke.c"), "-o", exe, "-L", libdir, "-lb200icp", "-lm",
           "-Wl,-rpath," + libdir]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr
    return exe


def test_header_is_plain_c_and_a_c_program_links(tmp_path):
    """include/b200icp.h compiles as C99 with -Wall -Werror, a C caller links against the library,
    and without a device the call fails loudly (exit code 2 = the library's error string path)."""
    import subprocess
    exe = _build_c_caller(tmp_path)
    if not torch.cuda.is_available():
        out = subprocess.run([exe], capture_output=True, text=True, timeout=120)
        assert out.returncode == 2 and "no sm_100" in out.stdout, (out.returncode, out.stdout)
