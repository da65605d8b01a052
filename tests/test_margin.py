"""Edge cases of the fp32 slack bound |d32 - d64| <= slack(d32, A) (csrc/icp_common.cuh::slack), beyond the
random clouds of tests/test_host.py::test_fp32_slack_bounds_the_fp64_distance: near-ties (two targets whose
float64 distances differ in the last bits), offsets of hundreds of metres (the cancellation the bound is
for), tiny and denormal separations, and points that collide after rounding to fp32.  The candidate set
{t : d32(t) <= m + 4 slack(m)} around the fp32 minimum m must contain the float64 nearest neighbour."""
import numpy as np
import pytest


def _slack(d_hi, a_bound):
    u = np.float32(2.0 ** -24)
    s = np.sqrt(np.float32(d_hi))
    a_s = np.float32(a_bound) + s
    return np.float32(8) * u * a_s * s + np.float32(8) * u * np.float32(d_hi) + np.float32(8) * u * u * a_s * a_s + np.float32(1.1754944e-38)


def _d32(q32, t32):
    diff = q32[None, :] - t32
    d = np.zeros(len(t32), np.float32)
    for a in range(t32.shape[1]):
        d = d + diff[:, a] * diff[:, a]
    return d


def _check(q, t):
    q32, t32 = q.astype(np.float32), t.astype(np.float32)
    d32 = _d32(q32, t32)
    d64 = ((q[None, :] - t) ** 2).sum(1)
    a_bound = np.float32(np.abs(q32).max() + np.abs(t32).max())
    with np.errstate(under="ignore"):
        assert np.all(np.abs(d32.astype(np.float64) - d64) <= _slack(d32, a_bound).astype(np.float64))
        m = d32.min()
        cand = d32 <= m + np.float32(4) * _slack(m, a_bound)
    assert cand[int(np.argmin(d64))], "the float64 nearest neighbour fell outside the candidate set"
    return int(cand.sum())


@pytest.mark.parametrize("offset", [0.0, 30.0, 750.0, 1.0e4])
@pytest.mark.parametrize("dim", [2, 3])
def test_near_ties_at_large_offsets(offset, dim):
    rng = np.random.default_rng(int(offset) + dim)
    for _ in range(200):
        q = rng.uniform(-1, 1, dim) + offset
        direction = rng.normal(size=dim); direction /= np.linalg.norm(direction)
        r = 10.0 ** rng.uniform(-4, 0)
        t0 = q + r * direction
        # a ring of targets at almost the same distance: relative differences from 1e-16 to 1e-6
        others = []
        for k in range(15):
            d2 = rng.normal(size=dim); d2 /= np.linalg.norm(d2)
            others.append(q + r * (1.0 + 10.0 ** rng.uniform(-16, -6)) * d2)
        _check(q, np.vstack([t0] + others))


@pytest.mark.parametrize("dim", [2, 3])
def test_tiny_and_denormal_separations(dim):
    rng = np.random.default_rng(5 + dim)
    for scale in (1e-3, 1e-12, 1e-20, 1e-30, 1e-42):        # squares underflow in fp32 from 1e-23 down
        for _ in range(50):
            q = rng.uniform(-1, 1, dim) * scale
            t = rng.uniform(-1, 1, (16, dim)) * scale
            _check(q, t)
    # targets that collide after rounding to fp32 but differ in float64
    base = np.full(dim, 12.3456789)
    t = np.vstack([base + k * 1e-9 for k in range(8)])
    assert len(np.unique(t.astype(np.float32), axis=0)) == 1
    n = _check(base + 3.2e-9, t)
    assert n == 8                                            # all of them are candidates; float64 decides


def test_query_on_top_of_a_target():
    t = np.array([[5.0, -3.0], [5.0, -3.0 + 1e-7], [5.0 + 1e-3, -3.0]])
    _check(t[0].copy(), t)
    _check(t[1].copy(), t)
