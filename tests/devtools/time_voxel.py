"""Developer timing of voxel_downsample on the device against the oracle (row F2)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
from lidar_slam_b200 import engine, synth
from oracle import icp_oracle as orc

for n in (25000, 200000, 2000000):
    pts = synth.drill_cloud(n, seed=1)
    d = torch.from_numpy(pts).cuda()
    for _ in range(2):
        out = engine.voxel_downsample_device(d, 0.005)
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(5):
        out = engine.voxel_downsample_device(d, 0.005)
    e.record(); e.synchronize()
    ms = s.elapsed_time(e) / 5
    line = f"n {n}: {ms:.3f} ms -> {out.shape[0]} voxels, {n / ms * 1e-3:.1f} M points/s, {n * 24 / ms * 1e-6:.1f} GB/s of input"
    if n <= 25000:
        t0 = time.time(); ref = orc.voxel_downsample(pts, 0.005); t1 = time.time()
        line += f"  (cpu oracle {1e3 * (t1 - t0):.0f} ms, equal: {np.array_equal(out.cpu().numpy(), ref)})"
    print(line)
