"""CPU: the occupancy-grid oracle against the reference's own known-answer vectors
(code/pr2_utils.py:94,96) and against whole-map outputs of the unmodified modules/ogm.py
(tests/golden/ogm_golden.npz, made by tests/golden/make_ogm_golden.py)."""
import os

import numpy as np
import pytest

from oracle import ogm_oracle as oo

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def ogm_golden():
    return np.load(os.path.join(ROOT, "tests", "golden", "ogm_golden.npz"), allow_pickle=False)


def split(points, counts):
    return np.split(points, np.cumsum(counts)[:-1])


def test_bresenham_known_answers_of_the_reference():
    # code/pr2_utils.py:93-96
    r1 = oo.bresenham2D(0, 1, 10, 5)
    assert np.array_equal(r1.T, [[0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10], [1, 1, 2, 2, 3, 3, 3, 4, 4, 5, 5]])
    r2 = oo.bresenham2D(0, 1, 9, 6)
    assert np.array_equal(r2.T, [[0, 1, 2, 3, 4, 5, 6, 7, 8, 9], [1, 2, 2, 3, 3, 4, 4, 5, 5, 6]])


def test_bresenham_closed_form_used_by_the_kernel():
    """The kernel addresses cell k directly: minor offset = (k*dy + dx - 1 - dx//2) // dx."""
    rng = np.random.default_rng(0)
    for _ in range(300):
        sx, sy, ex, ey = rng.integers(-40, 40, size=4)
        pts = oo.bresenham2D(sx, sy, ex, ey)
        dx, dy = abs(ex - sx), abs(ey - sy)
        steep = dy > dx
        if steep:
            dx, dy = dy, dx
        k = np.arange(dx + 1)
        m = (k * dy + dx - 1 - dx // 2) // dx if dx > 0 else np.zeros(1, dtype=int)
        if steep:
            want = np.stack([sx + np.sign(ex - sx) * m, sy + np.sign(ey - sy) * k], 1)
        else:
            want = np.stack([sx + np.sign(ex - sx) * k, sy + np.sign(ey - sy) * m], 1)
        assert np.array_equal(pts, want), (sx, sy, ex, ey)
        assert (pts[0] == (sx, sy)).all() and (pts[-1] == (ex, ey)).all()


def test_oracle_replays_reference_maps_bitwise(ogm_golden):
    g = ogm_golden
    for name in map(str, g["names"]):
        res, mxx, mxy, mnx, mny = g[f"{name}/params"]
        o = oo.OgmOracle(res, mxx, mxy, mnx, mny)
        o.build_map(g[f"{name}/poses"], split(g[f"{name}/points"], g[f"{name}/counts"]))
        assert o.log_odds.dtype == np.float32 and o.log_odds.shape == g[f"{name}/log_odds"].shape
        assert np.array_equal(o.log_odds, g[f"{name}/log_odds"])         # bit-exact float32
        assert np.array_equal(o.grid, g[f"{name}/grid"])


def test_binarisation_threshold_search_matches_reference_expression():
    from lidar_slam_b200 import ogm
    t = ogm.one_at_or_below()
    assert t < 0
    around = np.nextafter(np.float32(t), np.float32([-1, 1]))
    lo = np.array([around[0], t, around[1], 0.0, -0.0, 1e-7, -1e-7, -3e-7], dtype=np.float32)
    want = (1. / (1 + np.exp(lo))) > 0.5
    assert np.array_equal(lo <= t, want)
