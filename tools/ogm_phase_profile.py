import sys, os, ctypes
sys.path.insert(0, "/root/repo")
import numpy as np, torch
from lidar_slam_b200 import synth, ogm, _lib
from lidar_slam_b200.engine import _ptr, _stream_ptr
ns = 1000
poses = synth.make_trajectories(1, ns, seed=2)[0]
pool, counts = synth.cast_scans(poses, seed=2, device="cuda")
lib = _lib.load()
W, H = 361, 241
lo = torch.zeros((1, W, H), dtype=torch.float32, device="cuda")
nb = ctypes.c_size_t(); lib.b200icp_ogm_workspace_bytes(W, H, pool.shape[1], 1, ctypes.byref(nb))
ws = torch.zeros((nb.value,), dtype=torch.uint8, device="cuda")
cs = torch.as_tensor(np.stack([np.cos(poses[:, 2]), np.sin(poses[:, 2])], -1)).cuda()
pd = torch.as_tensor(poses).cuda()
for rep in range(2):
    lo.zero_()
    rc = lib.b200icp_ogm_build(_ptr(lo), W, H, 0.05, -9.0, -6.0, _ptr(pd), _ptr(cs), _ptr(pool), _ptr(counts), ns, pool.shape[1], 1, _ptr(ws), ws.numel(), _stream_ptr())
    torch.cuda.synchronize()
ctl = ws.view(torch.int32)[W * H: W * H + 16].cpu().numpy()
print("rc", rc, "us per phase total", ctl[4:10], "per scan", ctl[4:10] / ns)
print("phases: 0 ends+bbox, 1 walk/count, 2 reset+sweep, 3 walk/events, 4 mixed replay")
