"""Developer timing probe (not the benchmark): producer / consumer kernels either side of the ICP
path at config-2 size -- scan construction, occupancy-grid mapping, texture mapping -- with the
CPU oracle timed on a bounded sample beside each."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
from lidar_slam_b200 import synth, scans, ogm, texture_mapping as tm
from oracle import scan_oracle as so, ogm_oracle as oo, texture_oracle as to

def ev_time(fn, reps=3):
    best = 1e30
    for _ in range(reps):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); fn(); e.record(); e.synchronize()
        best = min(best, s.elapsed_time(e))
    return best

out = {}
n = int(sys.argv[1]) if len(sys.argv) > 1 else 4956
# ---- scan construction -------------------------------------------------------------------
gen = torch.Generator().manual_seed(1)
r = (4.0 + 3.0 * torch.rand((n, 1081), generator=gen, dtype=torch.float64))
r[torch.rand(r.shape, generator=gen) < 0.02] = 0.0
rd = r.cuda()
ms = ev_time(lambda: scans.scans_from_ranges(rd, 0.1, 30.0))
pool, counts = scans.scans_from_ranges(rd, 0.1, 30.0)
ms_f = ev_time(lambda: scans.statistical_outliers_removal_pool(pool.clone(), counts.clone(), 3))
t0 = time.perf_counter(); z = so.get_lidar_data(r[:500].numpy(), 0.1, 30.0); t1 = time.perf_counter()
so.statistical_outliers_removal(z, 3); t2 = time.perf_counter()
bytes_alg = n * 1081 * 8 + n * 1082 * 16
out["scan_build"] = {"scans": n, "gpu_ms": ms, "scans_per_s": n / ms * 1e3, "algorithmic_gbs": bytes_alg / ms / 1e6,
                     "filter_gpu_ms_incl_clone": ms_f, "cpu_oracle_scans_per_s": 500 / (t1 - t0),
                     "cpu_oracle_filter_scans_per_s": 500 / (t2 - t1)}
# ---- occupancy grid ------------------------------------------------------------------------
ns = min(n, 2000)
poses = synth.make_trajectories(1, ns, seed=2)[0]
pool, counts = synth.cast_scans(poses, seed=2, device="cuda")
for n_maps in (1, 3):
    lo = torch.zeros((n_maps, 361, 241), dtype=torch.float32, device="cuda")
    P = np.repeat(poses[None], n_maps, axis=0)
    ms = ev_time(lambda: ogm.build_maps(lo.zero_(), 0.05, -9.0, -6.0, P, (pool, counts)), reps=2)
    out[f"ogm_{n_maps}map"] = {"scans": ns, "rays_per_scan": float(counts.float().mean()), "gpu_ms": ms,
                               "scans_per_s_per_map": ns / ms * 1e3, "us_per_scan": ms * 1e3 / ns}
o = oo.OgmOracle(0.05, 9.0, 6.0, -9.0, -6.0)
zz = [pool[i, :counts[i]].cpu().numpy() for i in range(4)]
t0 = time.perf_counter(); o.build_map(poses[:4], zz); t1 = time.perf_counter()
out["ogm_cpu_oracle"] = {"scans_per_s": 4 / (t1 - t0)}
# ---- texture -------------------------------------------------------------------------------
F = 64
disp, rgb = to.synthetic_frames(F, 480, 640, seed=5)
x_ts = np.stack([np.linspace(-1, 1, F), np.linspace(0.5, -0.5, F), np.linspace(0, 3, F)], 1)
p = tm.TexturePainter(401, 401, 0.05, -10.0, -10.0, to.T_RC, to.K)
d_dev, c_dev = torch.from_numpy(disp).cuda(), torch.from_numpy(rgb).cuda()
ms = ev_time(lambda: p.paint_frames(d_dev, c_dev, x_ts))
grid = np.zeros((401, 401), np.uint8)
t0 = time.perf_counter()
to.generate_texture_map(x_ts[:4], np.arange(4.), np.arange(4.), np.arange(4.), disp[:4], rgb[:4], grid, 0.05, -10., -10., to.T_RC, to.K)
t1 = time.perf_counter()
out["texture"] = {"frames": F, "gpu_ms_incl_host_setup": ms, "frames_per_s": F / ms * 1e3,
                  "algorithmic_gbs": F * 480 * 640 * 7 / ms / 1e6, "cpu_oracle_frames_per_s": 4 / (t1 - t0)}
print(json.dumps(out, indent=1))
