"""The general (3-D) kernel, icp3d_flow.cuh: Morton-sorted block pruning against the same kernel with
every block swept (bit for bit), batch invariance of a pair's result (the summation order is a function
of the pair alone, not of the batch shape), the tie rule (lowest ORIGINAL index) under the Morton
permutation, the yaw-sweep call shape of code/icp_warm_up/test_icp.py:70-91 against the oracle, and
pair lists longer than one launch (1024 pairs)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _sweep_problem(n_src, n_targets, n_yaw=24, seed=1):
    import lidar_slam_b200 as eng
    from lidar_slam_b200 import synth
    src = synth.drill_cloud(n_src, seed=seed)
    tgts = [synth.partial_view(src, seed=10 * seed + k) for k in range(n_targets)]
    pool, counts = eng.pack_scans([src] + tgts)
    pairs, inits = [], []
    for j, t in enumerate(tgts):
        for T0 in synth.yaw_sweep_inits(src, t, n_yaw):
            pairs.append((0, 1 + j))
            inits.append(T0)
    return src, tgts, pool, counts, np.array(pairs, np.int32), np.stack(inits)


def _same(a, b):
    for name in ("T", "error", "iters", "status", "corr"):
        x, y = getattr(a, name), getattr(b, name)
        if x is not None:
            assert torch.equal(x, y), name


@pytest.mark.parametrize("n_src", [300, 2500, 7000])
def test_pruned_equals_unpruned_bitwise_and_oracle(n_src):
    import lidar_slam_b200 as eng
    from lidar_slam_b200 import engine
    from oracle import icp_oracle as orc
    src, tgts, pool, counts, pairs, inits = _sweep_problem(n_src, 2, n_yaw=6, seed=3)
    kw = dict(epsilon=1e-3, normalize_error=True, return_correspondences=True)
    a = eng.run_icp_batch(pool, counts, pairs, inits, **kw)
    b = eng.run_icp_batch(pool, counts, pairs, inits, flags=engine.FLAG_NO_PRUNE, **kw)
    _same(a, b)
    pruned_tiles = engine.swept_tiles()          # of the unpruned run
    eng.run_icp_batch(pool, counts, pairs, inits, **kw)
    assert engine.swept_tiles() < pruned_tiles or n_src < 1000
    for k in (0, 4, 7, 11):
        tgt = tgts[pairs[k, 1] - 1]
        To, eo, co, po = orc.run_icp_info(src, tgt, inits[k], epsilon=1e-3, normalize_error=True)
        assert int(a.iters[k]) == po
        assert np.abs(a.T[k].cpu().numpy() - To).max() < 1e-8
        assert float(a.error[k]) == pytest.approx(eo, rel=1e-9)
        assert np.array_equal(a.corr[k, :len(src)].cpu().numpy(), co)


def test_result_of_a_pair_does_not_depend_on_the_batch():
    """The same (source, target, init) alone, in a batch of 24 and in a batch of 96 with other clouds:
    identical bits (round 1's cluster kernel summed in an order chosen from the batch shape)."""
    import lidar_slam_b200 as eng
    src, tgts, pool, counts, pairs, inits = _sweep_problem(3000, 4, seed=5)
    kw = dict(epsilon=1e-3, normalize_error=True, return_correspondences=True)
    full = eng.run_icp_batch(pool, counts, pairs, inits, **kw)
    for k in (0, 29, 77):
        alone = eng.run_icp_batch(pool, counts, pairs[k:k + 1], inits[k:k + 1], **kw)
        assert torch.equal(alone.T[0], full.T[k]) and torch.equal(alone.error[0], full.error[k])
        assert torch.equal(alone.corr[0], full.corr[k]) and int(alone.iters[0]) == int(full.iters[k])
    part = eng.run_icp_batch(pool, counts, pairs[24:48], inits[24:48], **kw)
    assert torch.equal(part.T, full.T[24:48]) and torch.equal(part.iters, full.iters[24:48])
    perm = np.random.default_rng(0).permutation(len(pairs))
    shuf = eng.run_icp_batch(pool, counts, pairs[perm], inits[perm], **kw)
    assert torch.equal(shuf.T, full.T[perm]) and torch.equal(shuf.error, full.error[perm])


def test_exact_ties_go_to_the_lowest_original_index():
    """Duplicate target points far apart in index: the Morton sort puts them side by side in some
    order; the match must still be the LOWEST original index (the brute-force specification)."""
    import lidar_slam_b200 as eng
    from lidar_slam_b200 import engine
    from oracle import icp_oracle as orc
    rng = np.random.default_rng(4)
    tgt = rng.uniform(-1, 1, size=(900, 3))
    dup = rng.choice(900, size=120, replace=False)
    tgt = np.concatenate([tgt, tgt[dup]])                     # originals at < 900, copies at >= 900
    tgt = tgt[rng.permutation(len(tgt))]
    src = tgt[rng.choice(len(tgt), size=500)] + rng.normal(scale=1e-3, size=(500, 3))
    pool, counts = eng.pack_scans([src, tgt])
    corr, d2 = engine.nearest_batch(pool, counts, np.array([[0, 1]], np.int32), return_dist2=True)
    want = orc.nearest_brute(src, tgt) if hasattr(orc, "nearest_brute") else \
        ((src[:, None, :] - tgt[None, :, :]) ** 2).sum(-1).argmin(1)
    assert np.array_equal(corr[0, :500].cpu().numpy(), want)


def test_more_pairs_than_one_launch_and_big_planar_scans():
    """2300 pairs (three launches of <= 1024) of small clouds, and planar scans above the 2-D kernel's
    2048-point limit (dim 2 through the general kernel)."""
    import lidar_slam_b200 as eng
    from oracle import icp_oracle as orc
    rng = np.random.default_rng(8)
    clouds, pairs, inits = [], [], []
    for k in range(50):
        t = rng.uniform(-1, 1, size=(rng.integers(40, 90), 3))
        clouds.append(t)
    for k in range(2300):
        a, b = rng.integers(0, 50, size=2)
        pairs.append((a, b))
    pool, counts = eng.pack_scans(clouds)
    pairs = np.array(pairs, np.int32)
    res = eng.run_icp_batch(pool, counts, pairs, None, epsilon=1e-6, max_iters=15)
    assert bool((res.status <= 2).all())
    for k in (0, 1023, 1024, 2047, 2048, 2299):
        To, eo, _, po = orc.run_icp_info(clouds[pairs[k, 0]], clouds[pairs[k, 1]], np.eye(4), epsilon=1e-6,
                                         max_iters=15, nn="brute")
        assert int(res.iters[k]) == po and np.abs(res.T[k].cpu().numpy() - To).max() < 1e-8, k
    ang = np.linspace(-2.3, 2.3, 2600)
    wall = np.stack([3.0 * np.cos(ang) + 0.3 * np.cos(7 * ang), 2.0 * np.sin(ang)], axis=1)
    th = 0.02
    Rm = np.array([[np.cos(th), -np.sin(th)], [np.sin(th), np.cos(th)]])
    src = wall[::1] @ Rm.T + np.array([0.02, -0.01]) + rng.normal(scale=1e-3, size=wall.shape)
    pool2, counts2 = eng.pack_scans([src, wall])
    r2 = eng.run_icp_batch(pool2, counts2, np.array([[0, 1]], np.int32), None, epsilon=1e-4, return_correspondences=True)
    To, eo, co, po = orc.run_icp_info(src, wall, np.eye(4), epsilon=1e-4)
    assert int(r2.iters[0]) == po and np.abs(r2.T[0].cpu().numpy() - To).max() < 1e-8
    assert np.array_equal(r2.corr[0, :len(src)].cpu().numpy(), co)
