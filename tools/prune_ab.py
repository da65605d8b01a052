"""A/B of the block-pruned and the exhaustive sweep (developer tool).
   python tools/prune_ab.py run OUT.npz [n_pairs]   # one variant (env B200ICP_SWEEP), results + timing
   python tools/prune_ab.py cmp A.npz B.npz         # bit-for-bit comparison of two runs"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np


def run(out, n):
    import torch
    from lidar_slam_b200 import engine, synth
    peak = engine.ffma_peak_tflops()
    res_all = {}
    for name, b, kw in (
            ("odo", synth.odometry_batch(n, seed=20, device="cuda", n_traj=1), dict(epsilon=1e-3)),
            ("odo2", synth.odometry_batch(min(n, 1500), seed=77, device="cuda"), dict(epsilon=1e-3, normalize_error=True))):
        ms_best = 1e30
        for rep in range(4):
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            res = engine.run_icp_batch(b["scans"], b["counts"], b["pairs"], b["init"], return_correspondences=(rep == 0),
                                       planar_fallback=False, **kw)
            e.record(); e.synchronize()
            if rep == 0:
                first = res
            else:
                ms_best = min(ms_best, s.elapsed_time(e))
        tiles = engine.swept_tiles()
        it = res.iters.double(); cn = b["counts"].double(); pr = b["pairs"].long()
        cand = float((it * cn[pr[:, 0]] * cn[pr[:, 1]]).sum().item())
        npairs = b["pairs"].shape[0]
        print(f"{name}: sweep={os.environ.get('B200ICP_SWEEP', 'pruned')} {ms_best:.3f} ms {npairs / ms_best * 1e3:.0f} pairs/s "
              f"mean passes {it.mean().item():.2f}; algorithmic frac {cand * 4 / ms_best * 1e3 / peak['lane_ffma_per_s']:.3f}; "
              f"evaluated/algorithmic {tiles * 1024 / cand:.3f}; executed frac {tiles * 1024 * 4 / ms_best * 1e3 / peak['lane_ffma_per_s']:.3f}")
        for k in ("T", "error", "iters", "status", "corr"):
            res_all[name + "_" + k] = getattr(first, k).cpu().numpy()
    np.savez(out, **res_all)


def cmp(a, b):
    A, B = np.load(a), np.load(b)
    bad = 0
    for k in A.files:
        same = np.array_equal(A[k], B[k], equal_nan=A[k].dtype.kind == "f")
        print(k, "identical" if same else "DIFFERENT", A[k].shape)
        bad += not same
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    if sys.argv[1] == "run":
        run(sys.argv[2], int(sys.argv[3]) if len(sys.argv) > 3 else 4955)
    else:
        cmp(sys.argv[2], sys.argv[3])
