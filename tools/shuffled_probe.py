"""Developer probe: the pruned and the exhaustive sweep on scans whose points are in random order
(blocks of 32 consecutive points are then not spatially coherent, so hardly anything can be skipped)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from lidar_slam_b200 import engine, synth

n = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
b = synth.odometry_batch(n, seed=20, device="cuda", n_traj=1)
scans, counts = b["scans"].clone(), b["counts"]
g = torch.Generator(device="cuda").manual_seed(1)
for i in range(scans.shape[0]):
    c = int(counts[i])
    scans[i, :c] = scans[i, :c][torch.randperm(c, device="cuda", generator=g)]
for name, pool in (("ordered", b["scans"]), ("shuffled", scans)):
    for sweep in ("pruned", "exhaustive"):
        if sweep == "exhaustive":
            os.environ["B200ICP_SWEEP"] = "exhaustive"
        else:
            os.environ.pop("B200ICP_SWEEP", None)
        best = 1e9
        for rep in range(3):
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            res = engine.run_icp_batch(pool, counts, b["pairs"], b["init"], epsilon=1e-3, planar_fallback=False)
            e.record(); e.synchronize()
            if rep:
                best = min(best, s.elapsed_time(e))
        tiles = engine.swept_tiles()
        it = res.iters.double(); cn = counts.double(); pr = b["pairs"].long()
        cand = float((it * cn[pr[:, 0]] * cn[pr[:, 1]]).sum().item())
        print(f"{name:9s} {sweep:10s}: {best:8.3f} ms  {n / best * 1e3:9.0f} pairs/s  evaluated share {tiles * 1024 / cand:.3f}  "
              f"passes {it.mean().item():.2f}")
os.environ.pop("B200ICP_SWEEP", None)
