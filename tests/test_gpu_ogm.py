"""GPU parity of occupancy-grid mapping through the C ABI: float32 log-odds
grids and uint8 occupancy grids bit-identical to the reference's (golden vectors) and to the
oracle on fresh seeded inputs; batched maps; properties at full size."""
import os

import numpy as np
import pytest
import torch

from oracle import ogm_oracle as oo

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def ogm_golden():
    return np.load(os.path.join(ROOT, "tests", "golden", "ogm_golden.npz"), allow_pickle=False)


def split(points, counts):
    return np.split(points, np.cumsum(counts)[:-1])


def test_maps_bitwise_vs_reference_golden(ogm_golden):
    from lidar_slam_b200.ogm import OccupancyGridMap
    g = ogm_golden
    for name in map(str, g["names"]):
        res, mxx, mxy, mnx, mny = g[f"{name}/params"]
        m = OccupancyGridMap(res, mxx, mxy, mnx, mny)
        m.build_map(g[f"{name}/poses"], split(g[f"{name}/points"], g[f"{name}/counts"]))
        lo = m.grid_map_log_odds
        assert lo.dtype == np.float32 and lo.shape == g[f"{name}/log_odds"].shape
        bad = np.argwhere(lo != g[f"{name}/log_odds"])
        assert bad.shape[0] == 0, (name, bad[:5], lo[tuple(bad[0])], g[f"{name}/log_odds"][tuple(bad[0])])
        assert np.array_equal(m.grid_map, g[f"{name}/grid"])


def test_update_map_incremental_equals_build_map(ogm_golden):
    from lidar_slam_b200.ogm import OccupancyGridMap
    g = ogm_golden
    res, mxx, mxy, mnx, mny = g["clipped/params"]
    scans = split(g["clipped/points"], g["clipped/counts"])
    m = OccupancyGridMap(res, mxx, mxy, mnx, mny)
    for x_t, z_t in zip(g["clipped/poses"], scans):
        m.update_map(x_t, z_t)
    assert np.array_equal(m.grid_map_log_odds, g["clipped/log_odds"])


def test_batched_maps_and_oracle_on_fresh_inputs():
    """Three pose sequences over the same scans in one launch (the maps share every launch); each equals the
    oracle run alone.  Includes an empty scan and rays that start outside the grid."""
    from lidar_slam_b200 import ogm, synth
    n = 10
    poses = synth.make_trajectories(3, n, seed=9)
    poses[..., :2] *= 0.6
    scans, counts = synth.cast_scans(poses[0], seed=9)
    scans, counts = scans.numpy(), counts.numpy()
    meas = [scans[i, :counts[i]][::6] for i in range(n)]
    meas[4] = np.zeros((0, 2))
    poses[2, :, 0] += 7.0                                         # third trajectory leaves the small grid
    res, box = 0.1, (6.0, 4.0, -6.0, -4.0)
    ref = []
    for k in range(3):
        o = oo.OgmOracle(res, *box)
        o.build_map(poses[k], meas)
        ref.append(o)
    lo = torch.zeros((3, ref[0].width, ref[0].height), dtype=torch.float32, device="cuda")
    ogm.build_maps(lo, res, box[2], box[3], poses, meas)
    got = lo.cpu().numpy()
    for k in range(3):
        assert np.array_equal(got[k], ref[k].log_odds), k
        assert np.array_equal(ogm.binarize(lo[k]).cpu().numpy(), ref[k].grid)


def test_full_size_properties():
    """1081-beam scans on a 0.05 m grid: untouched cells stay 0, values stay in [-20, 20], every
    ray end cell inside the grid was hit, and rebuilding is deterministic."""
    from lidar_slam_b200 import ogm, synth
    n = 60
    poses = synth.make_trajectories(1, n, seed=2)[0]
    pool, counts = synth.cast_scans(poses, seed=2, device="cuda")
    m = ogm.OccupancyGridMap(0.05, 9.0, 6.0, -9.0, -6.0)
    m.build_map(poses, (pool, counts))
    a = m.grid_map_log_odds
    m2 = ogm.OccupancyGridMap(0.05, 9.0, 6.0, -9.0, -6.0)
    m2.build_map(poses, (pool, counts))
    assert np.array_equal(a, m2.grid_map_log_odds)
    assert a.min() >= -20 and a.max() <= 20 and (a != 0).sum() > 10000
    o = oo.OgmOracle(0.05, 9.0, 6.0, -9.0, -6.0)                     # spot-check the first two scans
    mm = ogm.OccupancyGridMap(0.05, 9.0, 6.0, -9.0, -6.0)
    z = [pool[i, :counts[i]].cpu().numpy() for i in range(2)]
    o.build_map(poses[:2], z)
    mm.build_map(poses[:2], z)
    assert np.array_equal(mm.grid_map_log_odds, o.log_odds) and np.array_equal(mm.grid_map, o.grid)


def test_event_list_overflow_path_is_exact(ogm_golden, monkeypatch):
    """With no room for event lists every cell that sees both signs in one scan is replayed by the
    closed-form ray test; with room for only a few lists both paths mix.  Same bits either way."""
    from lidar_slam_b200.ogm import OccupancyGridMap
    g = ogm_golden
    for per_ray in ("0", "1"):
        monkeypatch.setenv("B200ICP_OGM_EVENTS_PER_RAY", per_ray)
        for name in map(str, g["names"]):
            res, mxx, mxy, mnx, mny = g[f"{name}/params"]
            m = OccupancyGridMap(res, mxx, mxy, mnx, mny)
            m.build_map(g[f"{name}/poses"], split(g[f"{name}/points"], g[f"{name}/counts"]))
            assert np.array_equal(m.grid_map_log_odds, g[f"{name}/log_odds"]), (per_ray, name)


def test_many_scans_cross_batch_boundaries():
    """More scans than one batch holds (64): the planes are reused batch after batch; equals the
    oracle on every cell."""
    from lidar_slam_b200 import ogm, synth
    n = 150
    poses = synth.make_trajectories(1, n, seed=4)[0]
    poses[:, :2] *= 0.5
    scans, counts = synth.cast_scans(poses, seed=4)
    scans, counts = scans.numpy(), counts.numpy()
    meas = [scans[i, :counts[i]][::9] for i in range(n)]
    res, box = 0.1, (6.0, 4.0, -6.0, -4.0)
    o = oo.OgmOracle(res, *box)
    o.build_map(poses, meas)
    m = ogm.OccupancyGridMap(res, *box)
    m.build_map(poses, meas)
    assert np.array_equal(m.grid_map_log_odds, o.log_odds)
    assert np.array_equal(m.grid_map, o.grid)


@pytest.mark.parametrize("short_every", [2, 12])
def test_cells_with_long_mixed_event_lists(short_every):
    """A dense fan of rays where every `short_every`-th ray stops on a near arc that the others
    cross: the near-arc cells collect hundreds of interleaved decrements and increments per scan
    (long event lists: the gap replay for few increments, the general sort for many), some scans
    from close by so that the lists also meet cells far from the clip."""
    from lidar_slam_b200 import ogm
    rng = np.random.default_rng(short_every)
    n_scans, n_rays = 12, 900
    ang = np.linspace(-0.35, 0.35, n_rays)
    meas, poses = [], []
    for s in range(n_scans):
        near = 0.9 + 0.05 * rng.random(n_rays)
        far = 2.4 + 0.3 * rng.random(n_rays)
        r = np.where(np.arange(n_rays) % short_every == (s % short_every), near, far)
        meas.append(np.stack([r * np.cos(ang), r * np.sin(ang)], axis=1))
        poses.append([-1.0 + 0.02 * s, 0.03 * s, 0.1 * np.sin(s)])
    poses = np.array(poses)
    res, box = 0.1, (3.0, 2.0, -3.0, -2.0)
    o = oo.OgmOracle(res, *box)
    o.build_map(poses, meas)
    m = ogm.OccupancyGridMap(res, *box)
    m.build_map(poses, meas)
    assert np.array_equal(m.grid_map_log_odds, o.log_odds)
    assert np.array_equal(m.grid_map, o.grid)


def test_fine_grid_uses_short_batches():
    """540 k cells: the count planes of one batch are capped at 48 MB, so a batch holds 23 scans
    and 60 scans take three batches over two buffer sets; equals the oracle on every cell."""
    from lidar_slam_b200 import ogm, synth
    n = 60
    poses = synth.make_trajectories(1, n, seed=11)[0]
    scans, counts = synth.cast_scans(poses, seed=11)
    scans, counts = scans.numpy(), counts.numpy()
    meas = [scans[i, :counts[i]][::5] for i in range(n)]
    res, box = 0.02, (9.0, 6.0, -9.0, -6.0)
    o = oo.OgmOracle(res, *box)
    o.build_map(poses, meas)
    assert o.width * o.height * 4 * 64 > (48 << 20)
    m = ogm.OccupancyGridMap(res, *box)
    m.build_map(poses, meas)
    assert np.array_equal(m.grid_map_log_odds, o.log_odds)
    assert np.array_equal(m.grid_map, o.grid)
