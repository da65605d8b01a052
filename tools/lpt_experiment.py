import sys, os
sys.path.insert(0, "/root/repo")
import torch, numpy as np
import lidar_slam_b200 as eng
from lidar_slam_b200 import synth
b = synth.odometry_batch(4955, seed=20, device="cuda")
def run(pairs, init):
    best = 1e9
    for _ in range(4):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); res = eng.run_icp_batch(b["scans"], b["counts"], pairs, init, epsilon=1e-3); e.record(); e.synchronize()
        best = min(best, s.elapsed_time(e))
    return best, res
t0, res = run(b["pairs"], b["init"])
it = res.iters
order = torch.argsort(it, descending=True)
t1, _ = run(b["pairs"][order], b["init"][order])
order2 = torch.argsort(it, descending=False)
t2, _ = run(b["pairs"][order2], b["init"][order2])
# first-pass error as predictor
r1 = eng.run_icp_batch(b["scans"], b["counts"], b["pairs"], b["init"], epsilon=1e-3, max_iters=0)
pred = torch.argsort(r1.error, descending=True)
t3, _ = run(b["pairs"][pred], b["init"][pred])
print(f"natural {t0:.3f} ms; LPT by true passes {t1:.3f}; SPT {t2:.3f}; by first-pass error {t3:.3f}")
print("corr(first-pass err, passes)", np.corrcoef(r1.error.cpu().numpy(), it.cpu().numpy().astype(float))[0,1])
h = torch.bincount(it)
print("passes histogram (count per pass number):", h.tolist())
