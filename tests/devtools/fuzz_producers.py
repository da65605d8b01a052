"""Randomised bit-exactness run of scan construction, the range filter and the texture splat against
their oracles (developer tool, GPU box).  python tests/devtools/fuzz_producers.py [seconds] [seed]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
from lidar_slam_b200 import scans, texture_mapping as tm
from oracle import scan_oracle as so, texture_oracle as to

budget = float(sys.argv[1]) if len(sys.argv) > 1 else 60.0
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 1)
t_end = time.time() + budget
n_scan_cases = n_tex_cases = 0
while time.time() < t_end:
    # ---- scan construction + statistical range filter -------------------------------------
    n_scans, n_beams = int(rng.integers(1, 60)), int(rng.choice([1081, 1081, 361, 37, 2]))
    r = rng.uniform(0.0, 35.0, size=(n_scans, n_beams))
    r[rng.random(r.shape) < 0.05] = np.nan
    r[rng.random(r.shape) < 0.02] = rng.choice([0.1, 30.0])           # exactly on the gate
    r = np.float32(r).astype(np.float64) if rng.random() < 0.5 else r
    want = so.get_lidar_data(r, 0.1, 30.0)
    got = scans.get_lidar_data(r, 0.1, 30.0)
    assert len(want) == len(got)
    for a, b in zip(want, got):
        assert a.shape == b.shape and np.array_equal(a, b), "scan construction differs"
    if sum(len(z) for z in want) > 0:
        k_std = float(rng.choice([1, 2, 3]))
        wf, gf = so.statistical_outliers_removal(want, k_std), scans.statistical_outliers_removal(got, k_std)
        for a, b in zip(wf, gf):
            assert a.shape == b.shape and np.array_equal(a, b), "range filter differs"
    n_scan_cases += 1
    # ---- texture splat -----------------------------------------------------------------------
    F, h, w = int(rng.integers(1, 6)), int(rng.choice([24, 60])), int(rng.choice([32, 80]))
    disp, rgb = to.synthetic_frames(F, h, w, seed=int(rng.integers(1 << 30)))
    res = float(rng.choice([0.05, 0.1]))
    hx, hy = rng.uniform(1.0, 5.0), rng.uniform(1.0, 4.0)
    W, H = int(np.ceil(2 * hx / res + 1)), int(np.ceil(2 * hy / res + 1))
    grid = rng.integers(0, 2, size=(W, H)).astype(np.uint8)
    x_ts = np.stack([rng.uniform(-hx, hx, F), rng.uniform(-hy, hy, F), rng.uniform(-3.2, 3.2, F)], 1)
    stamps = np.arange(F, dtype=np.float64)
    _, want_tex = to.generate_texture_map(x_ts, stamps, stamps, stamps, disp, rgb, grid, res, -hx, -hy, to.T_RC, to.K)
    p = tm.TexturePainter(W, H, res, -hx, -hy, to.T_RC, to.K)
    p.paint_frames(disp, rgb, x_ts)
    got_tex = p.resolve(grid).cpu().numpy()
    assert np.array_equal(got_tex, want_tex), "texture differs"
    n_tex_cases += 1
print(f"producer fuzz ok: {n_scan_cases} scan-construction/filter cases, {n_tex_cases} texture cases, all bit-identical")
