import sys, time; sys.path.insert(0, '/root/repo')
import numpy as np, torch
from lidar_slam_b200 import ogm, synth
n = 300
poses = synth.make_trajectories(1, n, seed=2)[0]
pool, counts = synth.cast_scans(poses, seed=2)
z = [pool[i, :counts[i]].numpy() for i in range(n)]
m = ogm.OccupancyGridMap(0.05, 9.0, 6.0, -9.0, -6.0)
for i in range(20): m.update_map(poses[i], z[i])
torch.cuda.synchronize(); t = time.time()
for i in range(20, n): m.update_map(poses[i], z[i])
torch.cuda.synchronize(); print("update_map per call: %.1f us" % (1e6 * (time.time() - t) / (n - 20)))
