"""Latency of the un-batched drop-in: run_icp called pair after pair like modules/localization.py:112-130."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from lidar_slam_b200 import synth, hostapi
import lidar_slam_b200.icp as icp

b = synth.odometry_batch(400, seed=20)
b = {k: (v.numpy() if hasattr(v, "numpy") else v) for k, v in b.items()}
clouds = [b["scans"][i, :b["counts"][i]] for i in range(b["scans"].shape[0])]
for name, fn in (("icp.run_icp (torch engine)", lambda s, t, T0: icp.run_icp(s, t, init_transform=T0, epsilon=1e-3)),):
    for k in range(20):
        fn(clouds[b["pairs"][k][0]], clouds[b["pairs"][k][1]], b["init"][k])
    torch.cuda.synchronize(); t0 = time.time()
    for k in range(20, 400):
        fn(clouds[b["pairs"][k][0]], clouds[b["pairs"][k][1]], b["init"][k])
    torch.cuda.synchronize(); dt = time.time() - t0
    print(f"{name}: {1e6 * dt / 380:.0f} us per call")
# the host entry, one pair per call
t0 = None
for k in range(400):
    if k == 20:
        t0 = time.time()
    s, t = b["pairs"][k]
    pool, counts = hostapi.pack_scans_numpy([clouds[s], clouds[t]])
    hostapi.run_icp_batch_host(pool, counts, np.array([[0, 1]], np.int32), b["init"][k:k + 1], epsilon=1e-3)
print(f"hostapi one pair per call (incl. packing): {1e6 * (time.time() - t0) / 380:.0f} us per call")
