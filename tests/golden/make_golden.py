"""Generate golden input/output vectors by running the UNMODIFIED reference.

Run in the authoring container only (needs /root/reference):

    python tests/golden/make_golden.py

It imports /root/reference/modules/icp.py as-is, feeds it seeded synthetic
inputs from lidar_slam_b200.synth and stores inputs + the reference's outputs
(pose, error, correspondences, number of executed passes) in
tests/golden/icp_golden.npz.  The pass count is observed by wrapping the
reference's own `icp_iteration` with a counter; nothing else is touched.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.environ.get("LIDAR_SLAM_REF", "/root/reference"))

import modules.icp as ref  # noqa: E402  (the reference, unmodified)
from lidar_slam_b200 import synth  # noqa: E402

_calls = [0]
_orig_iteration = ref.icp_iteration


def _counting_iteration(*a, **k):
    _calls[0] += 1
    return _orig_iteration(*a, **k)


ref.icp_iteration = _counting_iteration


def run(src, tgt, init, eps, max_iters, stop, normalize):
    _calls[0] = 0
    T, err, corr = ref.run_icp(src, tgt, init_transform=init, epsilon=eps, max_iters=max_iters,
                               stopping_thresh=stop, return_error=True,
                               normalize_error=normalize, return_correspondences=True)
    return T, float(err), np.asarray(corr, dtype=np.int64), _calls[0]


def main():
    out = {}
    cases = []  # (name, src_key, tgt_key, init, eps, max_iters, stop, normalize)

    # ---- 2-D consecutive pairs (localization.py:120-123 call shape) ----------
    b = synth.odometry_batch(13, seed=7)
    scans, counts = b["scans"].numpy(), b["counts"].numpy()
    init = b["init"].numpy()
    for i in range(scans.shape[0]):
        out[f"scan{i}"] = scans[i, :counts[i]]
    for i in range(8):
        cases.append((f"odom{i}", f"scan{i+1}", f"scan{i}", init[i], 1e-3, 2000, 1e-4, False))
    # ---- loop-closure call shape (main.py:162): identity init, normalised ----
    for i in range(3):
        cases.append((f"loop{i}", f"scan{i}", f"scan{i+10}", np.eye(4), 1e-2, 2000, 1e-4, True))
    # same but forced to iterate (epsilon tiny) so normalised error drives the stop test
    cases.append(("loop_iter", "scan0", "scan10", np.eye(4), 1e-9, 2000, 1e-7, True))
    # ---- ragged, thinned scans (config 3 shape), N_s != N_t -------------------
    ts, tc = synth.thin_scans(b["scans"][:6], b["counts"][:6], seed=3)
    ts, tc = ts.numpy(), tc.numpy()
    for i in range(6):
        out[f"thin{i}"] = ts[i, :tc[i]]
    for i in range(4):
        cases.append((f"ragged{i}", f"thin{i+1}", f"thin{i}", init[i], 1e-3, 2000, 1e-4, False))
    # ---- max_iters exhaustion -------------------------------------------------
    cases.append(("maxit0", "scan3", "scan2", init[2], 1e-9, 0, 0.0, False))
    cases.append(("maxit3", "scan3", "scan2", init[2], 1e-9, 3, 0.0, False))
    # ---- reflection branch: target is the mirror image of the source ----------
    rng = np.random.default_rng(11)
    blob = rng.normal(size=(300, 2)) * np.array([2.0, 0.7])
    out["blob"] = blob
    out["blob_mirror"] = blob * np.array([1.0, -1.0]) + np.array([0.3, -0.2]) \
        + rng.normal(scale=1e-3, size=blob.shape)
    cases.append(("reflect", "blob", "blob_mirror", np.eye(4), 1e-6, 50, 1e-9, False))
    # ---- tiny clouds -----------------------------------------------------------
    out["tiny_a"] = rng.normal(size=(5, 2))
    out["tiny_b"] = out["tiny_a"][[2, 0, 4, 1]] + 0.01
    cases.append(("tiny", "tiny_a", "tiny_b", np.eye(4), 1e-6, 20, 1e-9, False))
    # ---- 3-D yaw sweep (test_icp.py:70-91 call shape) ---------------------------
    drill = synth.drill_cloud(1500, seed=5)
    view = synth.partial_view(drill, seed=6)
    out["drill"], out["drill_view"] = drill, view
    sweep = synth.yaw_sweep_inits(drill, view, 24)
    for k in (0, 5, 10, 15, 20, 23):
        cases.append((f"drill_yaw{k}", "drill", "drill_view", sweep[k], 1e-3, 2000, 1e-4, True))
    cont = synth.container_cloud(900, seed=8)
    cview = synth.partial_view(cont, seed=9)
    out["cont"], out["cont_view"] = cont, cview
    csweep = synth.yaw_sweep_inits(cont, cview, 24)
    for k in (3, 14):
        cases.append((f"cont_yaw{k}", "cont", "cont_view", csweep[k], 1e-3, 2000, 1e-4, True))
    # 3-D raw SSE + general (non-planar) initial transform
    th = 0.3
    tilt = np.eye(4)
    tilt[:3, :3] = np.array([[1, 0, 0], [0, np.cos(th), -np.sin(th)], [0, np.sin(th), np.cos(th)]])
    cases.append(("drill_tilt", "drill", "drill_view", sweep[2] @ tilt, 1e-3, 30, 1e-6, False))

    names = []
    for name, sk, tk, T0, eps, mi, st, nz in cases:
        T, err, corr, passes = run(out[sk], out[tk], T0, eps, mi, st, nz)
        names.append(name)
        out[f"case/{name}/src"] = np.array(sk)
        out[f"case/{name}/tgt"] = np.array(tk)
        out[f"case/{name}/init"] = np.asarray(T0, dtype=np.float64)
        out[f"case/{name}/params"] = np.array([eps, mi, st, float(nz)])
        out[f"case/{name}/T"] = T
        out[f"case/{name}/err"] = np.array(err)
        out[f"case/{name}/corr"] = corr
        out[f"case/{name}/passes"] = np.array(passes)
        print(f"{name:12s} passes={passes:3d} err={err:.6g}")
    out["case_names"] = np.array(names)

    # ---- voxel_downsample (icp.py:4-27) -----------------------------------------
    vox_in = synth.drill_cloud(3000, seed=12)
    out["voxel_in"] = vox_in
    out["voxel_size"] = np.array(0.01)
    out["voxel_out"] = ref.voxel_downsample(vox_in, 0.01)
    print("voxel", vox_in.shape, "->", out["voxel_out"].shape)

    np.savez_compressed(os.path.join(HERE, "icp_golden.npz"), **out)
    print("wrote", os.path.join(HERE, "icp_golden.npz"))


if __name__ == "__main__":
    main()
