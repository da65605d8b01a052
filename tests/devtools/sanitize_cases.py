"""Tiny run of every kernel for compute-sanitizer, ONE tool per gpurun call (memcheck, racecheck, synccheck):
   compute-sanitizer --tool memcheck python tests/devtools/sanitize_cases.py   -> profiles/r2_sanitizer_*.txt"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
import lidar_slam_b200 as eng
from lidar_slam_b200 import synth, ogm, scans, texture_mapping as tm, engine
from oracle import texture_oracle as to

b = synth.odometry_batch(24, seed=5, device="cuda")
r = eng.run_icp_batch(b["scans"], b["counts"], b["pairs"], b["init"], epsilon=1e-3, return_correspondences=True)
ts, tc = synth.thin_scans(b["scans"][:8].cpu(), b["counts"][:8].cpu(), seed=3)
eng.run_icp_batch(ts.cuda(), tc.cuda(), b["pairs"][:6], b["init"][:6], epsilon=1e-3, normalize_error=True)
engine.pack_poses(r.T, r.error)
src = synth.drill_cloud(1500, seed=1)
tgts = [synth.partial_view(src, seed=10 + k) for k in range(2)]
pool, counts = eng.pack_scans([src] + tgts)
pairs = torch.tensor([(0, 1), (0, 2), (0, 1)], dtype=torch.int32).cuda()
inits = torch.tensor(np.stack([synth.yaw_sweep_inits(src, tgts[0], 24)[k] for k in (0, 5, 9)])).cuda()
eng.run_icp_batch(pool, counts, pairs, inits, epsilon=1e-3, normalize_error=True, return_correspondences=True)
engine.nearest_batch(pool, counts, pairs[:2], return_dist2=True)
engine.voxel_downsample_device(torch.from_numpy(src).cuda(), 0.01)
rng = np.random.default_rng(0)
p2, c2 = scans.scans_from_ranges(rng.uniform(0, 32, size=(9, 1081)), 0.1, 30.0)
scans.statistical_outliers_removal_pool(p2, c2, 2)
poses = synth.make_trajectories(1, 6, seed=2)[0]
sp, sc = synth.cast_scans(poses, seed=2, device="cuda")
m = ogm.OccupancyGridMap(0.05, 4.0, 3.0, -4.0, -3.0)          # small grid: rays leave it
m.build_map(poses, (sp, sc))
disp, rgb = to.synthetic_frames(2, 120, 160, seed=5)
p = tm.TexturePainter(81, 61, 0.1, -4.0, -3.0, to.T_RC, to.K)
p.paint_frames(disp, rgb, poses[:2])
p.resolve(m.grid_map[:81, :61].copy())
# gated host entries: two slices, sorted and reversed pair lists; ranges in; the single-pair entry
from lidar_slam_b200 import hostapi, _lib
g = synth.odometry_batch(40, seed=6, device="cuda", with_ranges=True)
h = {k: g[k].cpu().numpy() for k in ("scans", "counts", "pairs", "init", "ranges")}
for pr, ini in ((h["pairs"], h["init"]), (h["pairs"][::-1].copy(), h["init"][::-1].copy())):
    hostapi.run_icp_batch_host(h["scans"], h["counts"], pr, ini, epsilon=1e-3, flags=_lib.flag_slices(2),
                               return_correspondences=True)
    hostapi.run_icp_ranges_host(h["ranges"], pr, ini, epsilon=1e-3, flags=_lib.flag_slices(2))
engine.run_icp_ranges_batch(g["ranges"], g["pairs"], g["init"], epsilon=1e-3, return_correspondences=True)
eng.run_icp(h["scans"][1, :h["counts"][1]], h["scans"][0, :h["counts"][0]], init_transform=h["init"][0], epsilon=1e-3,
            return_error=True, return_correspondences=True)
eng.run_icp_batch(b["scans"], b["counts"], b["pairs"], b["init"], epsilon=1e-3, flags=engine.FLAG_EXHAUSTIVE)
eng.run_icp_batch(b["scans"], b["counts"], b["pairs"], b["init"], epsilon=1e-3, flags=engine.FLAG_NO_PRUNE)
# the general kernel with several chunks per pass and shared targets
src2 = synth.drill_cloud(2600, seed=4)
tg2 = [synth.partial_view(src2, seed=20 + k) for k in range(2)]
pool2, counts2 = eng.pack_scans([src2] + tg2)
prs2 = torch.tensor([(0, 1), (0, 2), (0, 1), (0, 2)], dtype=torch.int32).cuda()
ini2 = torch.tensor(np.stack([synth.yaw_sweep_inits(src2, tg2[0], 24)[k] for k in (0, 3, 7, 11)])).cuda()
eng.run_icp_batch(pool2, counts2, prs2, ini2, epsilon=1e-3, normalize_error=True, return_correspondences=True)
# the allocation-free entry (scratch in the library's device image), general and planar kernel
import ctypes
lib = _lib.load()
vp = lambda t: ctypes.c_void_p(t.data_ptr())
for pool_t, counts_t, pairs_t, init_t, dim in ((pool2, counts2, prs2, ini2, 3), (b["scans"], b["counts"], b["pairs"], b["init"], 2)):
    n = int(pairs_t.shape[0])
    pairs_t = pairs_t.to("cuda", torch.int32).contiguous(); init_t = init_t.to("cuda", torch.float64).contiguous()
    eps_t = torch.full((n,), 1e-3, dtype=torch.float64, device="cuda"); nz = torch.zeros((n,), dtype=torch.uint8, device="cuda")
    T = torch.empty((n, 4, 4), dtype=torch.float64, device="cuda"); er = torch.empty((n,), dtype=torch.float64, device="cuda")
    it = torch.empty((n,), dtype=torch.int32, device="cuda"); stt = torch.empty((n,), dtype=torch.int32, device="cuda")
    rc = lib.b200icp_run_static(vp(pool_t), vp(counts_t), int(pool_t.shape[0]), int(pool_t.shape[1]), dim, vp(pairs_t), vp(init_t),
                                vp(eps_t), vp(nz), n, 2000, 1e-4, vp(T), vp(er), vp(it), vp(stt), None, None)
    assert rc == 0, rc
torch.cuda.synchronize()
print("sanitize cases done")
