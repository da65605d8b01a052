"""Developer timing probe (not the benchmark): config-2 shaped batch on one GPU."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import lidar_slam_b200 as eng
from lidar_slam_b200 import synth, engine

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4955
t0 = time.time()
b = synth.odometry_batch(n, seed=20, device="cuda")
torch.cuda.synchronize(); print("gen s", time.time() - t0)
peak = engine.ffma_peak_tflops()
print("ffma peak", peak)
variants = sys.argv[2].split(",") if len(sys.argv) > 2 else [""]
for rep in range(4 * len(variants)):
    os.environ["B200ICP_VARIANT"] = variants[rep % len(variants)]
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    res = eng.run_icp_batch(b["scans"], b["counts"], b["pairs"], b["init"], epsilon=1e-3)
    e.record(); e.synchronize()
    ms = s.elapsed_time(e)
    it = res.iters.double()
    cn = b["counts"].double()
    pr = b["pairs"].long()
    cand = (it * cn[pr[:, 0]] * cn[pr[:, 1]]).sum().item()
    print(f"rep{rep} var={variants[rep % len(variants)]!r}: {ms:.3f} ms  {n/ms*1e3:.0f} pairs/s  mean iters {it.mean().item():.2f} max {it.max().item():.0f} "
          f"pair-iters/s {it.sum().item()/ms*1e3:.3e}  lane-instr/s {cand*4/ms*1e3:.3e}  frac_of_probe {cand*4/ms*1e3/peak['lane_ffma_per_s']:.3f}")
print(torch.bincount(res.status).tolist())
