"""GPU parity tests: the CUDA engine (through the C ABI) against the golden vectors the
reference produced and against the CPU oracle on the same seeded inputs.

Tolerances (BASELINE.json north_star): poses within 1e-5 m / 1e-5 rad of the reference's
float64 result; correspondences identical (no exact ties in these inputs).  In practice we
also require the SAME number of passes and agreement to ~1e-9, far inside the bar.
"""
import numpy as np
import pytest
import torch

from conftest import golden_cases

pytestmark = pytest.mark.gpu

POSE_TOL = 1e-5          # the bar
TIGHT = 1e-9             # what we actually expect


def _engine():
    import lidar_slam_b200 as eng
    return eng


def _pose_delta(A, B):
    rot = np.abs(A[:3, :3] - B[:3, :3]).max()      # ~ radians for small differences
    trans = np.abs(A[:3, 3] - B[:3, 3]).max()
    return rot, trans


def test_golden_through_dropin_run_icp(golden):
    eng = _engine()
    for c in golden_cases(golden):
        T, err, corr = eng.run_icp(c["src"], c["tgt"], init_transform=c["init"], epsilon=c["epsilon"],
                                   max_iters=c["max_iters"], stopping_thresh=c["stopping_thresh"],
                                   return_error=True, normalize_error=c["normalize"],
                                   return_correspondences=True)
        rot, trans = _pose_delta(T, c["T"])
        assert rot < POSE_TOL and trans < POSE_TOL, (c["name"], rot, trans)
        assert rot < TIGHT and trans < TIGHT, (c["name"], rot, trans)
        assert err == pytest.approx(c["err"], rel=1e-9), c["name"]
        assert corr.dtype == np.int64 and np.array_equal(corr, c["corr"]), c["name"]
        assert isinstance(err, np.float64) and T.dtype == np.float64 and T.shape == (4, 4)


def test_golden_pass_counts_batched(golden):
    """All golden cases of one dimensionality in ONE launch, mixed epsilon / normalise flags."""
    eng = _engine()
    cases = list(golden_cases(golden))
    for dim in (2, 3):
        sel = [c for c in cases if c["src"].shape[1] == dim and c["max_iters"] == 2000
               and c["stopping_thresh"] == 1e-4]
        clouds, pairs = [], []
        for c in sel:
            pairs.append([len(clouds), len(clouds) + 1])
            clouds += [c["src"], c["tgt"]]
        scans, counts = eng.pack_scans(clouds)
        res = eng.run_icp_batch(scans, counts, np.array(pairs, np.int32),
                                np.stack([c["init"] for c in sel]),
                                epsilon=np.array([c["epsilon"] for c in sel]),
                                normalize_error=np.array([c["normalize"] for c in sel]),
                                return_correspondences=True)
        iters = res.iters.cpu().numpy()
        T = res.T.cpu().numpy()
        err = res.error.cpu().numpy()
        corr = res.corr.cpu().numpy()
        for k, c in enumerate(sel):
            assert iters[k] == c["passes"], (c["name"], iters[k], c["passes"])
            rot, trans = _pose_delta(T[k], c["T"])
            assert rot < TIGHT and trans < TIGHT, c["name"]
            assert err[k] == pytest.approx(c["err"], rel=1e-9)
            n = c["src"].shape[0]
            assert np.array_equal(corr[k, :n], c["corr"]) and np.all(corr[k, n:] == -1)


def test_return_conventions_match_reference(golden):
    eng = _engine()
    c = next(golden_cases(golden))
    kw = dict(init_transform=c["init"], epsilon=c["epsilon"])
    T = eng.run_icp(c["src"], c["tgt"], **kw)
    assert isinstance(T, np.ndarray) and T.shape == (4, 4)
    T2, e = eng.run_icp(c["src"], c["tgt"], return_error=True, **kw)
    T3, k = eng.run_icp(c["src"], c["tgt"], return_correspondences=True, **kw)
    T4, e4, k4 = eng.run_icp(c["src"], c["tgt"], return_error=True, return_correspondences=True, **kw)
    assert np.array_equal(T, T2) and np.array_equal(T, T3) and np.array_equal(T, T4)
    assert e == e4 and np.array_equal(k, k4)
    src0, tgt0, init0 = c["src"].copy(), c["tgt"].copy(), c["init"].copy()
    eng.run_icp(c["src"], c["tgt"], **kw)
    assert np.array_equal(src0, c["src"]) and np.array_equal(tgt0, c["tgt"]) and np.array_equal(init0, c["init"])


def test_odometry_batch_against_oracle():
    """Config-2 shaped pairs: iteration counts, correspondences and poses equal the oracle."""
    eng = _engine()
    from lidar_slam_b200 import synth
    from oracle import icp_oracle as orc
    b = synth.odometry_batch(96, seed=31, device="cuda")
    res = eng.run_icp_batch(b["scans"], b["counts"], b["pairs"], b["init"], epsilon=1e-3,
                            return_correspondences=True)
    scans, counts = b["scans"].cpu().numpy(), b["counts"].cpu().numpy()
    pairs, init = b["pairs"].cpu().numpy(), b["init"].cpu().numpy()
    T, err = res.T.cpu().numpy(), res.error.cpu().numpy()
    iters, corr = res.iters.cpu().numpy(), res.corr.cpu().numpy()
    status = res.status.cpu().numpy()
    for k in range(96):
        s, t = pairs[k]
        src, tgt = scans[s, :counts[s]], scans[t, :counts[t]]
        To, eo, co, po = orc.run_icp_info(src, tgt, init[k], epsilon=1e-3)
        assert iters[k] == po, (k, iters[k], po)
        rot, trans = _pose_delta(T[k], To)
        assert rot < TIGHT and trans < TIGHT, (k, rot, trans)
        assert err[k] == pytest.approx(eo, rel=1e-9)
        assert np.array_equal(corr[k, :counts[s]], co), k
        assert status[k] in (0, 2)


def test_ragged_mixed_batch_against_oracle():
    """Config-3 shape: thinned ragged scans, odometry pairs (raw SSE, eps 1e-3) and
    loop-closure pairs (identity init, eps 1e-2, normalised) in one launch."""
    eng = _engine()
    from lidar_slam_b200 import synth
    from oracle import icp_oracle as orc
    b = synth.odometry_batch(40, seed=77)
    scans, counts = synth.thin_scans(b["scans"], b["counts"], seed=5)
    odo = b["pairs"].numpy()
    loops = np.array([[i, i + 10] for i in range(0, 30, 10)], np.int32)
    pairs = np.concatenate([odo, loops])
    init = np.concatenate([b["init"].numpy(), np.tile(np.eye(4), (len(loops), 1, 1))])
    eps = np.concatenate([np.full(len(odo), 1e-3), np.full(len(loops), 1e-2)])
    norm = np.concatenate([np.zeros(len(odo), bool), np.ones(len(loops), bool)])
    res = eng.run_icp_batch(scans.cuda(), counts.cuda(), pairs, init, epsilon=eps, normalize_error=norm,
                            return_correspondences=True)
    sc, cn = scans.numpy(), counts.numpy()
    T, err, iters = res.T.cpu().numpy(), res.error.cpu().numpy(), res.iters.cpu().numpy()
    corr = res.corr.cpu().numpy()
    for k in range(len(pairs)):
        s, t = pairs[k]
        To, eo, co, po = orc.run_icp_info(sc[s, :cn[s]], sc[t, :cn[t]], init[k], epsilon=eps[k],
                                          normalize_error=bool(norm[k]))
        assert iters[k] == po, k
        rot, trans = _pose_delta(T[k], To)
        assert rot < TIGHT and trans < TIGHT, k
        assert err[k] == pytest.approx(eo, rel=1e-9)
        assert np.array_equal(corr[k, :cn[s]], co)


@pytest.mark.parametrize("dim,n_src,n_tgt", [(2, 1081, 1081), (2, 200, 37), (2, 33, 2000), (2, 3000, 2500),
                                             (3, 1500, 800), (3, 5000, 4100), (3, 70, 9)])
def test_nearest_is_exact(dim, n_src, n_tgt):
    """Exact float64 nearest neighbour == exhaustive float64 search (oracle.nn_brute)."""
    eng = _engine()
    from oracle import icp_oracle as orc
    rng = np.random.default_rng(dim * 1000 + n_src)
    tgt = rng.normal(size=(n_tgt, dim)) * 3.0 + 5.0
    src = tgt[rng.integers(0, n_tgt, n_src)] + rng.normal(scale=0.02, size=(n_src, dim))
    scans, counts = eng.pack_scans([src, tgt])
    corr, d2 = eng.nearest_batch(scans, counts, np.array([[0, 1]], np.int32), return_dist2=True)
    got = corr[0, :n_src].cpu().numpy()
    want = orc.nn_brute(src, tgt)
    assert np.array_equal(got, want)
    np.testing.assert_allclose(d2[0, :n_src].cpu().numpy(), ((src - tgt[want]) ** 2).sum(1), rtol=1e-13)


def test_nearest_with_duplicates_and_near_ties():
    """Duplicate target points (exact ties) and targets that differ by 1e-9 m: distances must be
    the true minimum; on exact ties the lowest index is returned (the KD-tree's choice is
    arbitrary there)."""
    eng = _engine()
    from oracle import icp_oracle as orc
    rng = np.random.default_rng(3)
    base = rng.normal(size=(500, 2)) * 4.0
    tgt = np.concatenate([base, base[:100], base[100:200] + 1e-9])     # exact duplicates + near ties
    src = base + rng.normal(scale=1e-3, size=base.shape)
    scans, counts = eng.pack_scans([src, tgt])
    corr = eng.nearest_batch(scans, counts, np.array([[0, 1]], np.int32))[0, :500].cpu().numpy()
    want = orc.nn_brute(src, tgt)
    assert np.array_equal(corr, want)


def test_large_coordinates_keep_exactness():
    """Coordinates near the 30 m range limit plus a far outlier: slack bound must still hold."""
    eng = _engine()
    from oracle import icp_oracle as orc
    rng = np.random.default_rng(9)
    tgt = rng.uniform(-30, 30, size=(1081, 2))
    tgt[5] = [2.5e4, -1.0e4]
    src = tgt[rng.permutation(1081)] + rng.normal(scale=2e-4, size=(1081, 2))
    scans, counts = eng.pack_scans([src, tgt])
    corr = eng.nearest_batch(scans, counts, np.array([[0, 1]], np.int32))[0, :1081].cpu().numpy()
    assert np.array_equal(corr, orc.nn_brute(src, tgt))


def test_statuses_and_empty_and_skipped():
    eng = _engine()
    from lidar_slam_b200 import synth
    b = synth.odometry_batch(4, seed=2)
    counts = b["counts"].clone()
    counts[2] = 0                                     # scan 2 empty -> pairs touching it report EMPTY
    pairs = np.array([[1, 0], [2, 1], [3, 2], [-1, 0], [4, 3]], np.int32)
    init = np.concatenate([b["init"].numpy(), np.eye(4)[None]])[[0, 1, 2, 4, 3]]
    res = eng.run_icp_batch(b["scans"].cuda(), counts.cuda(), pairs, init, epsilon=1e-3)
    status = res.status.cpu().numpy()
    assert status[1] == 4 and status[2] == 4 and status[3] == 6
    assert status[0] in (0, 2) and status[4] in (0, 2)
    assert res.iters.cpu().numpy()[1] == 0
    # max_iters = 0 -> exactly one pass, status max_iters (icp.py:174)
    res = eng.run_icp_batch(b["scans"].cuda(), b["counts"].cuda(), pairs[:1], init[:1], epsilon=1e-12, max_iters=0)
    assert res.iters.item() == 1 and res.status.item() == 1
    with pytest.raises(ValueError):
        eng.run_icp(np.zeros((0, 2)), np.zeros((5, 2)))


def test_nonplanar_init_on_2d_scans_reroutes_to_general_kernel():
    eng = _engine()
    from lidar_slam_b200 import synth
    from oracle import icp_oracle as orc
    b = synth.odometry_batch(2, seed=4)
    s, c = b["scans"].numpy(), b["counts"].numpy()
    init = b["init"].numpy()[0].copy()
    init[2, 3] = 0.05                                 # lifts the scan out of the plane
    T, err = eng.run_icp(s[1, :c[1]], s[0, :c[0]], init_transform=init, epsilon=1e-3, return_error=True)
    To, eo, _, _ = orc.run_icp_info(s[1, :c[1]], s[0, :c[0]], init, epsilon=1e-3)
    rot, trans = _pose_delta(T, To)
    assert rot < TIGHT and trans < TIGHT and err == pytest.approx(eo, rel=1e-9)


def test_voxel_downsample_golden(golden):
    eng = _engine()
    out = eng.voxel_downsample(golden["voxel_in"], float(golden["voxel_size"]))
    assert out.shape == golden["voxel_out"].shape
    np.testing.assert_array_equal(out, golden["voxel_out"])


def test_voxel_downsample_large_against_oracle():
    eng = _engine()
    from lidar_slam_b200 import synth
    from oracle import icp_oracle as orc
    pts = synth.drill_cloud(25000, seed=1)
    out = eng.voxel_downsample(pts, 0.005)
    want = orc.voxel_downsample(pts, 0.005)
    assert out.shape == want.shape
    np.testing.assert_allclose(out, want, rtol=0, atol=1e-15)


def test_yaw_sweep_best_of_24_matches_oracle():
    """Config 1/4 shape: 24 yaw inits of one (source, target) in one launch; best error and
    pose equal the oracle's sweep (test_icp.py:70-93)."""
    eng = _engine()
    from lidar_slam_b200 import synth
    from oracle import icp_oracle as orc
    src = synth.drill_cloud(1200, seed=21)
    tgt = synth.partial_view(src, seed=22)
    inits = synth.yaw_sweep_inits(src, tgt, 24)
    scans, counts = eng.pack_scans([src, tgt])
    res = eng.run_icp_batch(scans, counts, np.tile(np.array([[0, 1]], np.int32), (24, 1)), inits,
                            epsilon=1e-3, normalize_error=True)
    err = res.error.cpu().numpy()
    iters = res.iters.cpu().numpy()
    T = res.T.cpu().numpy()
    for k in range(24):
        To, eo, _, po = orc.run_icp_info(src, tgt, inits[k], epsilon=1e-3, normalize_error=True)
        assert iters[k] == po, k
        assert err[k] == pytest.approx(eo, rel=1e-9)
        rot, trans = _pose_delta(T[k], To)
        assert rot < 1e-8 and trans < 1e-8, (k, rot, trans)
    assert round(float(err.min()), 3) == round(min(
        orc.run_icp_info(src, tgt, inits[k], epsilon=1e-3, normalize_error=True)[1] for k in range(24)), 3)


@pytest.mark.parametrize("n_src", [5, 31, 32, 33, 255, 257, 544, 1024, 1280, 2047, 2048, 2050])
def test_full_icp_across_row_layouts_and_kernel_boundaries(n_src):
    """Source sizes that exercise every row layout of the planar kernel (whole rows per warp, 0-7
    shared left-over rows, the <2,4>, <4,4>, <5,8> and <8,8> variants) and the hand-over to the
    general kernel above 2048 points: passes, poses, errors and matches equal the oracle."""
    eng = _engine()
    from oracle import icp_oracle as orc
    rng = np.random.default_rng(n_src)
    ang = np.sort(rng.uniform(-2.2, 2.2, n_src + 40))
    rad = 4.0 + 1.5 * np.sin(3 * ang) + 0.8 * np.sign(np.sin(7 * ang))
    tgt = np.stack([rad * np.cos(ang), rad * np.sin(ang)], 1) + rng.normal(scale=2e-3, size=(n_src + 40, 2))
    th, t = 0.012, np.array([0.02, -0.015])
    R = np.array([[np.cos(th), -np.sin(th)], [np.sin(th), np.cos(th)]])
    src = (tgt[20:20 + n_src] - t) @ R + rng.normal(scale=2e-3, size=(n_src, 2))
    T, err, corr = eng.run_icp(src, tgt, epsilon=1e-9, max_iters=40, return_error=True, return_correspondences=True)
    To, eo, co, po = orc.run_icp_info(src, tgt, np.eye(4), epsilon=1e-9, max_iters=40)
    rot, trans = _pose_delta(T, To)
    assert rot < TIGHT and trans < TIGHT, (rot, trans)
    assert err == pytest.approx(eo, rel=1e-9)
    assert np.array_equal(corr, co)


def test_reflection_branch_on_device_matches_oracle():
    """A mirrored source makes det(cross-covariance) < 0: the reference's SVD solution is then a
    REFLECTION of the plane (icp.py:62-67 on planar data); the device's closed
    form takes the same branch and reports R[2,2] = -1."""
    eng = _engine()
    from oracle import icp_oracle as orc
    rng = np.random.default_rng(8)
    n = 300
    s = rng.choice([-1.0, 1.0], n)
    tgt = np.stack([0.1 * np.arange(n), 0.01 * s * (1 + rng.random(n))], 1)      # a zig-zag along the x axis
    src = tgt * np.array([1.0, -1.0]) + rng.normal(scale=1e-4, size=tgt.shape)  # mirrored: point i still matches i
    T, err = eng.run_icp(src, tgt, epsilon=1e-12, max_iters=3, return_error=True)
    To, eo, _, po = orc.run_icp_info(src, tgt, np.eye(4), epsilon=1e-12, max_iters=3)
    assert To[2, 2] < 0                                # the oracle (== reference) did reflect
    assert np.abs(T - To).max() < 1e-9 and err == pytest.approx(eo, rel=1e-9)
