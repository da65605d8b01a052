"""Developer probe: the general kernel on growing clouds with host-side wall clocks and the control words."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
import lidar_slam_b200 as eng
from lidar_slam_b200 import synth, engine

sizes = [int(a) for a in sys.argv[1:]] or [8000, 12000, 16000, 20000]
for n_src in sizes:
    src = synth.drill_cloud(n_src, seed=1)
    tgts = [synth.partial_view(src, seed=10 + k) for k in range(4)]
    pool, counts = eng.pack_scans([src] + tgts)
    pairs, inits = [], []
    for j, t in enumerate(tgts):
        for T0 in synth.yaw_sweep_inits(src, t, 24):
            pairs.append((0, 1 + j)); inits.append(T0)
    pairs = torch.tensor(pairs, dtype=torch.int32).cuda()
    inits = torch.tensor(np.stack(inits)).cuda()
    for n_pairs in (1, 24, 96):
        for flags in (0, engine.FLAG_NO_PRUNE):
            if flags and n_pairs > 24:
                continue
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            res = eng.run_icp_batch(pool, counts, pairs[:n_pairs], inits[:n_pairs], epsilon=1e-3, normalize_error=True, flags=flags)
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
            ctl = engine._last_workspace[:32].cpu().numpy().view(np.uint32)
            it = res.iters.double()
            print(f"n_src {n_src} pairs {n_pairs} flags {flags}: {dt * 1e3:.1f} ms  passes mean {it.mean():.1f} max {it.max():.0f} "
                  f"status {sorted(set(res.status.cpu().tolist()))} ctl head {ctl[0]} tail {ctl[1]} done {ctl[4]} abort {ctl[5]} "
                  f"tiles {int(ctl[2]) + (int(ctl[3]) << 32)}", flush=True)
