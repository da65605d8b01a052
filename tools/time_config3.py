"""Developer timing of config 3 (main.py --mode gtsam --filter_lidar --fixed_interval 10 shape): ragged
filtered scans, 4955 odometry pairs (perturbed-odometry init, epsilon 1e-3, raw error) and 495
loop-closure pairs (i, i+10) with identity init, epsilon 1e-2 and normalised error (main.py:161-172)
in ONE launch; pruned against exhaustive sweep, results compared bit for bit."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from lidar_slam_b200 import engine, synth

n = 4955
b = synth.odometry_batch(n, seed=20, device="cpu", n_traj=1)
scans, counts = synth.thin_scans(b["scans"], b["counts"], seed=5)
loop_src = np.arange(0, n + 1 - 10, 10)[:495]
pairs = np.concatenate([b["pairs"].numpy(), np.stack([loop_src, loop_src + 10], axis=1).astype(np.int32)])
init = np.concatenate([b["init"].numpy(), np.tile(np.eye(4), (len(loop_src), 1, 1))])
eps = np.concatenate([np.full(n, 1e-3), np.full(len(loop_src), 1e-2)])
norm = np.concatenate([np.zeros(n, np.uint8), np.ones(len(loop_src), np.uint8)])
dev = lambda x: torch.as_tensor(x).cuda()
scans, counts, pairs_d, init_d, eps_d, norm_d = scans.cuda(), counts.cuda(), dev(pairs), dev(init), dev(eps), dev(norm)
out = {}
for sweep in ("pruned", "exhaustive"):
    if sweep == "exhaustive":
        os.environ["B200ICP_SWEEP"] = "exhaustive"
    best = 1e9
    for rep in range(4):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        res = engine.run_icp_batch(scans, counts, pairs_d, init_d, epsilon=eps_d, normalize_error=norm_d,
                                   planar_fallback=False)
        e.record(); e.synchronize()
        if rep:
            best = min(best, s.elapsed_time(e))
    out[sweep] = res
    it = res.iters.double()
    print(f"config 3, {sweep:10s}: {len(pairs)} pairs (mean {counts.double().mean().item():.0f} points per scan) in {best:.3f} ms = "
          f"{len(pairs) / best * 1e3:.0f} pairs/s; passes odometry {it[:n].mean().item():.2f}, loop {it[n:].mean().item():.2f}; "
          f"status counts {torch.bincount(res.status).tolist()}")
os.environ.pop("B200ICP_SWEEP", None)
a, e = out["pruned"], out["exhaustive"]
print("pruned == exhaustive:", bool(torch.equal(a.T, e.T) and torch.equal(a.error, e.error) and torch.equal(a.iters, e.iters)))
