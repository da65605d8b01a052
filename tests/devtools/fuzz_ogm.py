"""Randomised bit-exactness run of the occupancy grid against the oracle (developer tool, GPU box).
python tests/devtools/fuzz_ogm.py [seconds] [seed]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
from lidar_slam_b200 import ogm
from oracle import ogm_oracle as oo

budget = float(sys.argv[1]) if len(sys.argv) > 1 else 60.0
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 1)
t_end = time.time() + budget
n_cases = n_scans_total = 0
while time.time() < t_end:
    res = float(rng.choice([0.05, 0.1, 0.2, 0.37]))
    hx, hy = rng.uniform(2.0, 8.0), rng.uniform(2.0, 6.0)
    box = (hx, hy, -hx * rng.uniform(0.5, 1.0), -hy * rng.uniform(0.5, 1.0))
    n = int(rng.integers(1, 140))
    n_maps = int(rng.integers(1, 4))
    poses = np.zeros((n_maps, n, 3))
    for m in range(n_maps):
        poses[m, :, 0] = np.cumsum(rng.normal(scale=0.05, size=n)) + rng.uniform(-hx, hx) * 0.5
        poses[m, :, 1] = np.cumsum(rng.normal(scale=0.05, size=n)) + rng.uniform(-hy, hy) * 0.5
        poses[m, :, 2] = np.cumsum(rng.normal(scale=0.03, size=n)) + rng.uniform(-3, 3)
    meas = []
    for s in range(n):
        k = int(rng.integers(0, 160))
        ang = np.sort(rng.uniform(-2.35, 2.35, k))
        r = rng.choice([rng.uniform(0.1, 1.0), rng.uniform(1.0, 6.0), rng.uniform(6.0, 25.0)], size=k) * rng.uniform(0.9, 1.1, k)
        if rng.random() < 0.2 and k:
            r[:] = r[0]                                   # an arc: many rays through the same cells
        meas.append(np.stack([r * np.cos(ang), r * np.sin(ang)], 1).reshape(-1, 2))
    if max(len(z) for z in meas) == 0:
        continue
    ref = []
    for m in range(n_maps):
        o = oo.OgmOracle(res, *box)
        o.build_map(poses[m], meas)
        ref.append(o)
    lo = torch.zeros((n_maps, ref[0].width, ref[0].height), dtype=torch.float32, device="cuda")
    ogm.build_maps(lo, res, box[2], box[3], poses, meas)
    got = lo.cpu().numpy()
    for m in range(n_maps):
        if not np.array_equal(got[m], ref[m].log_odds):
            bad = np.argwhere(got[m] != ref[m].log_odds)
            print("MISMATCH case", n_cases, "map", m, "cells", bad[:5], got[m][tuple(bad[0])], ref[m].log_odds[tuple(bad[0])],
                  "res", res, "box", box, "n", n)
            sys.exit(1)
        if not np.array_equal(ogm.binarize(lo[m]).cpu().numpy(), ref[m].grid):
            print("MISMATCH grid", n_cases, m); sys.exit(1)
    n_cases += 1
    n_scans_total += n * n_maps
print(f"ogm fuzz ok: {n_cases} cases, {n_scans_total} map-scans, all cells bit-identical")
