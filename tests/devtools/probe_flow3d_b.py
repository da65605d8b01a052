"""Developer probe: nearest-neighbour pass and few-pass ICP of the general kernel on one big pair, against torch."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
import lidar_slam_b200 as eng
from lidar_slam_b200 import synth, engine

n_src = int(sys.argv[1]); mode = sys.argv[2]
src = synth.drill_cloud(n_src, seed=1)
tgt = synth.partial_view(src, seed=10)
pool, counts = eng.pack_scans([src, tgt])
pairs = torch.tensor([[0, 1]], dtype=torch.int32).cuda()
T0 = torch.tensor(synth.yaw_sweep_inits(src, tgt, 24)[:1]).cuda()
t0 = time.perf_counter()
if mode == "nn":
    corr, d2 = engine.nearest_batch(pool, counts, pairs, return_dist2=True)
    torch.cuda.synchronize()
    print("nearest ms", 1e3 * (time.perf_counter() - t0), flush=True)
    s, t = torch.tensor(src).cuda(), torch.tensor(tgt).cuda()
    want = torch.cdist(s, t).argmin(1)
    got = corr[0, :n_src].long()
    bad = (want != got).nonzero().flatten()
    print("mismatches", bad.numel(), "first", bad[:10].tolist(), got[bad[:10]].tolist(), want[bad[:10]].tolist(), flush=True)
else:
    flags = engine.FLAG_NO_PRUNE if mode == "noprune" else 0
    for mi in (1, 3, 8):
        t0 = time.perf_counter()
        res = eng.run_icp_batch(pool, counts, pairs, T0, epsilon=1e-3, normalize_error=True, max_iters=mi, flags=flags)
        torch.cuda.synchronize()
        ctl = engine._last_workspace[:32].cpu().numpy().view(np.uint32)
        print(f"n {n_src} mode {mode} max_iters {mi}: {1e3 * (time.perf_counter() - t0):.1f} ms iters {res.iters.tolist()} status {res.status.tolist()} "
              f"err {res.error.tolist()} ctl {ctl[:6].tolist()}", flush=True)
