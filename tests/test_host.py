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
    return np.float32(8) * u * a_s * s + np.float32(8) * u * np.float32(d_hi) + np.float32(8) * u * u * a_s * a_s


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
