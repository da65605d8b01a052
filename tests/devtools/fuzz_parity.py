"""Randomised parity run (developer tool, GPU box): random sizes / dimensions / motions / stop
parameters, device result against the oracle.  python tests/devtools/fuzz_parity.py [seconds] [seed]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
import lidar_slam_b200 as eng
from lidar_slam_b200 import hostapi
from oracle import icp_oracle as orc

budget = float(sys.argv[1]) if len(sys.argv) > 1 else 60.0
big = len(sys.argv) > 3 and sys.argv[3] == "big"      # clouds of 2100..6000 points: the general kernel, 2-D and 3-D
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 1)
t_end = time.time() + budget
n_cases = n_pairs = 0
worst = 0.0
while time.time() < t_end:
    dim = int(rng.choice([2, 2, 2, 3]))
    n_scans = int(rng.integers(2, 6))
    clouds = []
    base_n = int(rng.choice([rng.integers(24, 70), rng.integers(70, 600), rng.integers(600, 2300)]))   # tiny clouds: rank-deficient fits, see DESIGN.md section 2
    if big:
        base_n = int(rng.integers(2100, 6000))
    if dim == 2:
        ang = np.sort(rng.uniform(-2.3, 2.3, base_n + 50))
        rad = 3.0 + rng.uniform(0.5, 3.0) * np.sin(rng.integers(1, 6) * ang) + rng.uniform(0, 1.5) * np.sign(np.sin(9 * ang))
        world = np.stack([rad * np.cos(ang), rad * np.sin(ang)], 1)
    else:
        world = rng.normal(size=(base_n + 50, 3)) * np.array([1.0, 0.6, 0.3])
    for s in range(n_scans):
        keep = np.sort(rng.choice(world.shape[0], size=int(rng.integers(max(16, base_n // 2), base_n + 1)), replace=False))
        th = rng.normal(scale=0.02)
        t = rng.normal(scale=0.03, size=dim)
        R = np.eye(dim)
        R[:2, :2] = [[np.cos(th), -np.sin(th)], [np.sin(th), np.cos(th)]]
        clouds.append((world[keep] - t) @ R + rng.normal(scale=rng.choice([0.0, 1e-3, 5e-3]), size=(keep.size, dim)))
    pool, counts = hostapi.pack_scans_numpy(clouds)
    B = int(rng.integers(1, 9))
    pairs = rng.integers(0, n_scans, size=(B, 2)).astype(np.int32)
    eps = rng.choice([1e-2, 1e-3, 1e-6], size=B)
    norm = rng.random(B) < 0.4
    max_iters = int(rng.choice([3, 25, 60]))
    init = np.tile(np.eye(4), (B, 1, 1))
    for k in range(B):
        th = rng.normal(scale=0.01)
        init[k, :2, :2] = [[np.cos(th), -np.sin(th)], [np.sin(th), np.cos(th)]]
        init[k, :dim, 3] = rng.normal(scale=0.01, size=dim)
    res = hostapi.run_icp_batch_host(pool, counts, pairs, init, epsilon=eps, max_iters=max_iters,
                                     normalize_error=norm, return_correspondences=True)
    for k in range(B):
        s, t = pairs[k]
        To, eo, co, po = orc.run_icp_info(clouds[s], clouds[t], init[k], epsilon=eps[k], max_iters=max_iters,
                                          normalize_error=bool(norm[k]), nn="brute")
        d = float(np.abs(res.T[k] - To).max())
        worst = max(worst, d)
        ok = res.iters[k] == po and d < 1e-8 and np.array_equal(res.correspondences[k, :counts[s]], co) \
            and abs(res.error[k] - eo) <= 1e-9 * max(1.0, abs(eo))
        if not ok:
            np.savez("gpurun_out/fuzz_fail.npz", pool=pool, counts=counts, pairs=pairs, init=init, eps=eps, norm=norm,
                     max_iters=max_iters, k=k)
            print("MISMATCH case", n_cases, "pair", k, "dim", dim, "sizes", counts[s], counts[t], "iters", res.iters[k], po,
                  "dT", d, "err", res.error[k], eo, "corr equal", np.array_equal(res.correspondences[k, :counts[s]], co))
            sys.exit(1)
        n_pairs += 1
    n_cases += 1
print(f"fuzz ok: {n_cases} batches, {n_pairs} pairs, worst |T - T_oracle| = {worst:.2e}")
