"""Small, fixed workloads for ncu captures of the non-headline kernels (one case per run):
   python tests/devtools/profile_cases.py ogm|texture|icp3d|scans"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
import lidar_slam_b200 as eng
from lidar_slam_b200 import synth, ogm, scans, texture_mapping as tm
from oracle import texture_oracle as to

case = sys.argv[1]
if case == "ogm":
    ns = 200
    poses = synth.make_trajectories(1, ns, seed=2)[0]
    pool, counts = synth.cast_scans(poses, seed=2, device="cuda")
    lo = torch.zeros((1, 361, 241), dtype=torch.float32, device="cuda")
    for _ in range(2):
        ogm.build_maps(lo.zero_(), 0.05, -9.0, -6.0, poses[None], (pool, counts))
elif case == "texture":
    F = 32
    disp, rgb = to.synthetic_frames(F, 480, 640, seed=5)
    x_ts = np.stack([np.linspace(-1, 1, F), np.linspace(0.5, -0.5, F), np.linspace(0, 3, F)], 1)
    p = tm.TexturePainter(401, 401, 0.05, -10.0, -10.0, to.T_RC, to.K)
    d, c = torch.from_numpy(disp).cuda(), torch.from_numpy(rgb).cuda()
    for _ in range(2):
        p.paint_frames(d, c, x_ts)
    p.resolve(None)
elif case in ("icp3d", "icp3d_20k"):
    src = synth.drill_cloud(20000 if case == "icp3d_20k" else 6000, seed=1)
    tgts = [synth.partial_view(src, seed=10 + k) for k in range(4)]
    pool, counts = eng.pack_scans([src] + tgts)
    pairs, inits = [], []
    for j, t in enumerate(tgts):
        for T0 in synth.yaw_sweep_inits(src, t, 24):
            pairs.append((0, 1 + j)); inits.append(T0)
    pairs = torch.tensor(pairs, dtype=torch.int32).cuda()
    inits = torch.tensor(np.stack(inits)).cuda()
    for _ in range(2):
        eng.run_icp_batch(pool, counts, pairs, inits, epsilon=1e-3, normalize_error=True)
elif case == "scans":
    gen = torch.Generator().manual_seed(1)
    r = (4.0 + 3.0 * torch.rand((4956, 1081), generator=gen, dtype=torch.float64)).cuda()
    for _ in range(2):
        pool, counts = scans.scans_from_ranges(r, 0.1, 30.0)
        scans.statistical_outliers_removal_pool(pool, counts, 3)
torch.cuda.synchronize()
print("done", case)
