"""Ground work for the hybrid sweep: checks numerically that the block-centred
expanded form  v = (|t'|^2 - 2 q'.t') + |q'|^2  evaluated in float32 (two FMAs per candidate) stays within
E = 16 u (|q'|^2 + r_b^2) of the exact squared distance between the SAME float32 inputs.  FMA is emulated as
float64 multiply-add rounded once to float32 (products of float32 numbers are exact in float64).
python tools/expanded_bound_check.py [cases]"""
import sys
import numpy as np

u = 2.0 ** -24
f32 = np.float32


def fma(a, b, c):
    return (a.astype(np.float64) * b.astype(np.float64) + c.astype(np.float64)).astype(f32)


rng = np.random.default_rng(0)
cases = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
worst = 0.0
for _ in range(cases):
    r = 10.0 ** rng.uniform(-3, 1)                     # block radius 1 mm .. 10 m
    dq = 10.0 ** rng.uniform(-3, 1.5)                  # distance of the query from the block centre
    n = 64
    t = (rng.normal(size=(n, 2)) * r / 2).astype(f32)
    t = t[np.hypot(t[:, 0].astype(np.float64), t[:, 1].astype(np.float64)) <= r]
    if t.shape[0] == 0:
        continue
    ang = rng.uniform(0, 2 * np.pi)
    q = np.array([dq * np.cos(ang), dq * np.sin(ang)]).astype(f32)
    if rng.random() < 0.5:                              # a query right next to one of the targets
        q = (t[rng.integers(t.shape[0])].astype(np.float64) + rng.normal(scale=10.0 ** rng.uniform(-4, -1), size=2)).astype(f32)
    t2 = fma(t[:, 1], t[:, 1], (t[:, 0] * t[:, 0]).astype(f32))          # fl(ty^2 + fl(tx^2))
    a = (f32(-2.0) * q[0]).astype(f32)                                   # exact
    b = (f32(-2.0) * q[1]).astype(f32)
    w = fma(np.full_like(t2, b), t[:, 1], fma(np.full_like(t2, a), t[:, 0], t2))
    q2 = fma(q[1:2], q[1:2], (q[0:1] * q[0:1]).astype(f32))[0]
    v = (w + q2).astype(f32)
    exact = ((q[0].astype(np.float64) - t[:, 0].astype(np.float64)) ** 2 +
             (q[1].astype(np.float64) - t[:, 1].astype(np.float64)) ** 2)
    rb2 = float(np.max(t[:, 0].astype(np.float64) ** 2 + t[:, 1].astype(np.float64) ** 2))
    E = 16 * u * (float(q2) + rb2)
    ratio = float(np.max(np.abs(v.astype(np.float64) - exact)) / E)
    worst = max(worst, ratio)
    assert ratio <= 1.0, (r, dq, ratio)
print(f"expanded-form bound holds on {cases} random blocks; worst |v - exact| / E = {worst:.3f}")
