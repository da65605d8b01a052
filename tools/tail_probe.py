"""Developer probe: pass-count distribution of the config-2 batch and kernel time against batch size
(how much of the launch is the tail of the longest pairs)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from lidar_slam_b200 import engine, synth

for n in [int(a) for a in sys.argv[1:]] or [4955, 20000]:
    b = synth.odometry_batch(n, seed=20, device="cuda", n_traj=1)
    best = 1e9
    for rep in range(4):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        res = engine.run_icp_batch(b["scans"], b["counts"], b["pairs"], b["init"], epsilon=1e-3, planar_fallback=False)
        e.record(); e.synchronize()
        if rep:
            best = min(best, s.elapsed_time(e))
    it = res.iters.cpu().numpy()
    q = np.percentile(it, [50, 90, 99, 99.9, 100])
    print(f"n={n}: {best:.3f} ms, {n / best * 1e3:.0f} pairs/s; passes sum {it.sum()} mean {it.mean():.2f} "
          f"p50/p90/p99/p99.9/max {q.tolist()}; perfectly balanced over 444 CTAs the longest pair is "
          f"{it.max() / (it.sum() / 444):.2f} of a CTA's share")
