"""The block-pruned NN sweep of the planar kernel against its exhaustive variant and the oracle.

The shipped kernel (icp2d_rows.cuh) skips target blocks whose bounding-box lower bound exceeds the
candidate threshold; FLAG_NO_PRUNE sweeps every block in the same kernel and FLAG_EXHAUSTIVE runs the
register-tiled exhaustive kernel of round 1 (icp2d.cuh).  Pruned and unpruned must give the SAME bits
-- poses, errors, pass counts, correspondences -- on every input, including inputs on which pruning
has nothing to prune (shuffled clouds) and the degenerate block counts; the exhaustive kernel sums in
another order, so it must give the same matches, pass counts and statuses and poses equal to
rounding.  The oracle check (exhaustive float64 search) closes the triangle.
"""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _eng():
    import lidar_slam_b200 as eng
    return eng


def _both(fn):
    """fn(flags) under the shipped row kernel (pruned), the same kernel with the block bounds switched
    off (FLAG_NO_PRUNE: every block swept, same summation order) and the exhaustive register-tiled
    kernel of round 1 (FLAG_EXHAUSTIVE).  The first two must agree bit for bit; the third sums in
    another order, so its matches / pass counts / statuses must be equal and its poses and errors
    equal to rounding.  Returns (pruned, unpruned)."""
    from lidar_slam_b200 import engine
    a = fn(0)
    b = fn(engine.FLAG_NO_PRUNE)
    e = fn(engine.FLAG_EXHAUSTIVE)
    for name in ("iters", "status", "corr"):
        x, y = getattr(a, name), getattr(e, name)
        if x is not None:
            assert torch.equal(x, y), "exhaustive kernel: " + name
    fin = torch.isfinite(a.error)
    assert torch.equal(fin, torch.isfinite(e.error))
    assert torch.allclose(a.T[fin], e.T[fin], rtol=1e-11, atol=1e-11)
    assert torch.allclose(a.error[fin], e.error[fin], rtol=1e-11, atol=0)
    return a, b

def _same(a, b):
    for name in ("T", "error", "iters", "status", "corr"):
        x, y = getattr(a, name), getattr(b, name)
        if x is None:
            continue
        assert torch.equal(x, y) or (x.is_floating_point() and torch.equal(torch.nan_to_num(x, nan=-1.0),
                                                                           torch.nan_to_num(y, nan=-1.0))), name


def test_pruned_equals_exhaustive_on_lidar_batch():
    eng = _eng()
    from lidar_slam_b200 import engine, synth
    b = synth.odometry_batch(600, seed=3, device="cuda")
    run = lambda flags: eng.run_icp_batch(b["scans"], b["counts"], b["pairs"], b["init"], epsilon=1e-3,
                                          return_correspondences=True, flags=flags)
    a, e = _both(run)
    _same(a, e)
    # and pruning really happened on ordered scans (tile counter: workspace bytes 8..15)
    eng.run_icp_batch(b["scans"], b["counts"], b["pairs"], b["init"], epsilon=1e-3)
    pruned_tiles = engine.swept_tiles()
    eng.run_icp_batch(b["scans"], b["counts"], b["pairs"], b["init"], epsilon=1e-3, flags=engine.FLAG_NO_PRUNE)
    all_tiles = engine.swept_tiles()
    assert 0 < pruned_tiles < 0.25 * all_tiles


@pytest.mark.parametrize("n_src,n_tgt", [(1, 1), (5, 31), (40, 32), (33, 33), (200, 64), (256, 65), (300, 500),
                                         (512, 480), (700, 1081), (1081, 1081), (1280, 1200), (1500, 2048),
                                         (2048, 1999)])
def test_pruned_equals_exhaustive_and_oracle_on_shuffled_clouds(n_src, n_tgt):
    """Random, unordered clouds: blocks are not spatially coherent, the boxes overlap and (almost)
    nothing can be skipped.  Every CTA shape of the planar kernel (n_max buckets 256 / 512 / 1280 /
    2048) and the 1-, 2- and 64-block targets are covered."""
    eng = _eng()
    from oracle import icp_oracle as orc
    rng = np.random.default_rng(n_src * 7919 + n_tgt)
    tgt = rng.uniform(-6.0, 6.0, size=(n_tgt, 2))
    th = 0.05
    Rm = np.array([[np.cos(th), -np.sin(th)], [np.sin(th), np.cos(th)]])
    pick = rng.integers(0, n_tgt, size=n_src)
    src = (tgt[pick] - np.array([0.03, -0.02])) @ Rm + rng.normal(scale=0.01, size=(n_src, 2))
    scans, counts = eng.pack_scans([src, tgt])
    pairs = np.array([[0, 1]], np.int32)
    run = lambda flags: eng.run_icp_batch(scans, counts, pairs, None, epsilon=1e-6, max_iters=30,
                                          return_correspondences=True, flags=flags)
    a, e = _both(run)
    _same(a, e)
    T, err, corr, passes = orc.run_icp_info(src, tgt, np.eye(4), epsilon=1e-6, max_iters=30, nn="brute")
    assert int(a.iters[0]) == passes
    assert np.abs(a.T[0].cpu().numpy() - T).max() < 1e-9
    assert np.array_equal(a.corr[0, :n_src].cpu().numpy(), corr)


def test_pruned_equals_exhaustive_on_clustered_far_apart_targets():
    """Target blocks far from every query (skipped) next to blocks that tie with the best one
    (ambiguity fallback): duplicates across block borders, clusters metres apart."""
    eng = _eng()
    from oracle import icp_oracle as orc
    rng = np.random.default_rng(11)
    centres = rng.uniform(-25.0, 25.0, size=(12, 2))
    tgt = np.concatenate([c + rng.normal(scale=0.05, size=(48, 2)) for c in centres])
    tgt[31] = tgt[32]                       # an exact duplicate across the border of blocks 0 and 1
    tgt[95] = tgt[96] + 1e-9                # a near tie across blocks 2 and 3
    src = tgt[rng.permutation(len(tgt))[:400]] + rng.normal(scale=0.004, size=(400, 2))
    src[0] = tgt[31]
    src[1] = tgt[96]
    scans, counts = eng.pack_scans([src, tgt])
    pairs = np.array([[0, 1]], np.int32)
    run = lambda flags: eng.run_icp_batch(scans, counts, pairs, None, epsilon=1e-9, max_iters=12,
                                          return_correspondences=True, flags=flags)
    a, e = _both(run)
    _same(a, e)
    T, err, corr, passes = orc.run_icp_info(src, tgt, np.eye(4), epsilon=1e-9, max_iters=12, nn="brute")
    assert int(a.iters[0]) == passes
    assert np.array_equal(a.corr[0, :400].cpu().numpy(), corr)
    assert np.abs(a.T[0].cpu().numpy() - T).max() < 1e-9


def test_pruned_equals_exhaustive_with_large_offsets_and_normalised_error():
    """Coordinates of hundreds of metres (wide slack), normalised error, mixed epsilons, ragged scans."""
    eng = _eng()
    from lidar_slam_b200 import synth
    b = synth.odometry_batch(96, seed=9, device="cuda")
    scans = b["scans"].clone()
    scans[..., 0] += 512.25
    scans[..., 1] -= 777.5
    counts = b["counts"].clone()
    counts[::3] = counts[::3] // 2          # ragged
    counts[5] = 17
    eps = torch.full((96,), 1e-3, dtype=torch.float64)
    eps[::2] = 1e-2
    norm = torch.zeros(96, dtype=torch.uint8)
    norm[::4] = 1
    init = b["init"].clone()
    # the initial transforms were made for the unshifted scans: conjugate them by the shift
    shift = torch.eye(4, dtype=torch.float64, device=init.device)
    shift[0, 3], shift[1, 3] = 512.25, -777.5
    init = shift @ init @ torch.linalg.inv(shift)
    init[:, 2, :] = torch.tensor([0.0, 0.0, 1.0, 0.0], dtype=torch.float64)
    init[:, 3, :] = torch.tensor([0.0, 0.0, 0.0, 1.0], dtype=torch.float64)
    init[:, :2, 2] = 0.0
    run = lambda flags: eng.run_icp_batch(scans, counts, b["pairs"], init, epsilon=eps, normalize_error=norm,
                                          return_correspondences=True, flags=flags)
    a, e = _both(run)
    _same(a, e)
    assert bool(((a.status == 0) | (a.status == 2)).all())


def test_pruned_handles_non_finite_points_like_exhaustive():
    eng = _eng()
    rng = np.random.default_rng(2)
    tgt = rng.uniform(-3, 3, size=(100, 2))
    src = tgt[:64] + 0.01
    src[7, 0] = np.nan
    scans, counts = eng.pack_scans([src, tgt])
    from lidar_slam_b200 import engine
    run = lambda flags: eng.run_icp_batch(scans, counts, np.array([[0, 1]], np.int32), None, epsilon=1e-3,
                                          max_iters=5, flags=flags)
    a, e = run(0), run(engine.FLAG_EXHAUSTIVE)
    n = run(engine.FLAG_NO_PRUNE)
    assert int(n.status[0]) == 3 and int(n.iters[0]) == int(a.iters[0])
    assert int(a.status[0]) == int(e.status[0]) == 3       # non-finite error stops the pair
    assert int(a.iters[0]) == int(e.iters[0])
