"""Developer timing probe (not the benchmark): configs 1 / 4 -- the 24-yaw initialisation sweep of
code/icp_warm_up/test_icp.py on synthetic 3-D objects, all (target, yaw) problems in one launch."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
import lidar_slam_b200 as eng
from lidar_slam_b200 import synth, engine
from oracle import icp_oracle as orc

peak = engine.ffma_peak_tflops()["lane_ffma_per_s"]
for n_src, n_tgt_clouds in ((1000, 4), (6000, 4), (20000, 4)):
    src = synth.drill_cloud(n_src, seed=1)
    tgts = [synth.partial_view(src, seed=10 + k) for k in range(n_tgt_clouds)]
    clouds = [src] + tgts
    pool, counts = eng.pack_scans(clouds)
    pairs, inits = [], []
    for j, t in enumerate(tgts):
        for T0 in synth.yaw_sweep_inits(src, t, 24):
            pairs.append((0, 1 + j)); inits.append(T0)
    pairs = torch.tensor(pairs, dtype=torch.int32).cuda()
    inits = torch.tensor(np.stack(inits)).cuda()
    for rep in range(2):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        res = eng.run_icp_batch(pool, counts, pairs, inits, epsilon=1e-3, normalize_error=True)
        e.record(); e.synchronize()
        ms = s.elapsed_time(e)
    it = res.iters.double()
    cand = float((it * counts[pairs[:, 0].long()].double() * counts[pairs[:, 1].long()].double()).sum())
    t0 = time.perf_counter()
    orc.run_icp_info(src, tgts[0], inits[3].cpu().numpy(), epsilon=1e-3, normalize_error=True)
    cpu_s = time.perf_counter() - t0
    print(f"n_src {n_src} n_tgt ~{int(counts[1])} problems {len(pairs)}: {ms:.1f} ms  passes mean {it.mean():.1f} max {it.max():.0f} "
          f"pair-iterations/s {it.sum().item() / ms * 1e3:.3e}  6*cand/s / ffma_peak {6 * cand / (ms * 1e-3) / peak:.3f}  "
          f"(cpu oracle, one problem with {int(res.iters[3])} passes: {cpu_s * 1e3:.0f} ms)")
