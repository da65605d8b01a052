"""GPU parity of scan construction through the C ABI: bit-exact against the
reference's golden vectors and the oracle, plus size-independent properties at full size."""
import os

import numpy as np
import pytest
import torch

from oracle import scan_oracle as so

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def scan_golden():
    return np.load(os.path.join(ROOT, "tests", "golden", "scan_golden.npz"), allow_pickle=False)


def _split(points, counts):
    return np.split(points, np.cumsum(counts)[:-1])


def test_scans_from_ranges_bitwise_vs_reference_golden(scan_golden):
    from lidar_slam_b200 import scans
    g = scan_golden
    for name in map(str, g["names"]):
        pool, counts = scans.scans_from_ranges(g[f"{name}/ranges"], 0.1, 30.0)
        assert pool.shape[1] % 2 == 0 and pool.dtype == torch.float64
        assert np.array_equal(counts.cpu().numpy(), g[f"{name}/counts"])
        got = scans.pool_to_list(pool, counts)
        assert np.array_equal(np.vstack(got), g[f"{name}/points"])        # bit-exact
        p = pool.cpu().numpy()
        for i, c in enumerate(g[f"{name}/counts"]):
            assert not p[i, c:].any()                                      # tail zero-filled


def test_range_filter_vs_reference_golden(scan_golden):
    from lidar_slam_b200 import scans
    g = scan_golden
    for name in map(str, g["names"]):
        for k in (1, 2, 3):
            pool, counts = scans.scans_from_ranges(g[f"{name}/ranges"], 0.1, 30.0)
            z = scans.pool_to_list(pool, counts)
            pool, counts, stats = scans.statistical_outliers_removal_pool(pool, counts, k_std=k)
            mean, std, thr = so.range_statistics(z, k)
            st = stats.cpu().numpy()
            assert st[3] == sum(s.shape[0] for s in z)
            # the global sums are reduced in a different (fixed) order than numpy's pairwise sum
            assert abs(st[0] - mean) <= 4e-16 * mean and abs(st[1] - std) <= 1e-13 * std
            assert abs(st[2] - thr) <= 1e-13 * thr
            assert np.array_equal(counts.cpu().numpy(), g[f"{name}/k{k}/counts"])
            assert np.array_equal(np.vstack(scans.pool_to_list(pool, counts)), g[f"{name}/k{k}/points"])


def test_drop_in_list_functions_match_oracle():
    from lidar_slam_b200 import scans
    rng = np.random.default_rng(5)
    r = rng.uniform(0.0, 33.0, size=(40, 1081))
    r[rng.random(r.shape) < 0.3] = 0.0
    r[7] = 0.0                                                             # a scan with no valid beam
    got, want = scans.get_lidar_data(r, 0.1, 30.0), so.get_lidar_data(r, 0.1, 30.0)
    assert [a.shape for a in got] == [b.shape for b in want]
    assert all(np.array_equal(a, b) for a, b in zip(got, want))
    gf, wf = scans.statistical_outliers_removal(want, k_std=2), so.statistical_outliers_removal(want, k_std=2)
    assert [a.shape for a in gf] == [b.shape for b in wf]
    assert all(np.array_equal(a, b) for a, b in zip(gf, wf))


def test_full_size_properties():
    """Config-2 size (4956 scans x 1081 beams): counts equal the gate mask sums, filtering is
    idempotent at a fixed threshold ordering (second pass with k large keeps everything), and
    the pool feeds run_icp_batch directly."""
    from lidar_slam_b200 import scans
    import lidar_slam_b200 as eng
    gen = torch.Generator(device="cpu").manual_seed(3)
    r = (4.0 + torch.rand((4956, 1081), generator=gen, dtype=torch.float64) * 3.0)
    r[torch.rand(r.shape, generator=gen) < 0.02] = 0.0
    r = r.cuda()
    pool, counts = scans.scans_from_ranges(r, 0.1, 30.0)
    assert torch.equal(counts.long(), ((r >= 0.1) & (r <= 30.0)).sum(1))
    before = counts.clone()
    pool, counts, stats = scans.statistical_outliers_removal_pool(pool, counts, k_std=1e9)
    assert torch.equal(before, counts)
    pairs = torch.stack([torch.arange(1, 9), torch.arange(0, 8)], 1).int().cuda()
    res = eng.run_icp_batch(pool, counts, pairs, None, epsilon=1e-3, max_iters=3)
    assert bool(torch.isfinite(res.error).all())
