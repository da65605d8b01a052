"""TEST-ONLY stand-in for the device aligner, so that the multi-rank host logic
(lidar_slam_b200.engine.run_icp_sharded) can be exercised with gloo on a CPU box.
It routes through the oracle; it is never importable from the product package."""
import numpy as np
import torch

from oracle import icp_oracle as orc


def cpu_align(scans, counts, pairs, init, epsilon=0.01, max_iters=2000, stopping_thresh=1e-4,
              normalize_error=False):
    scans, counts = np.asarray(scans), np.asarray(counts)
    pairs, init = np.asarray(pairs).reshape(-1, 2), np.asarray(init).reshape(-1, 4, 4)
    out = np.zeros((len(pairs), 4))
    Ts = np.zeros((len(pairs), 4, 4))
    for k, (s, t) in enumerate(pairs):
        T, err, _, _ = orc.run_icp_info(scans[s, :counts[s]], scans[t, :counts[t]], init[k], epsilon,
                                        max_iters, stopping_thresh, normalize_error, nn="brute")
        Ts[k] = T
        out[k] = [T[0, 3], T[1, 3], np.arctan2(T[1, 0], T[0, 0]), err]
    return out


def install_cpu_standin(engine):
    def run_icp_batch(scans, counts, pairs, init_transforms=None, **kw):
        kw.pop("planar_fallback", None)
        kw.pop("return_correspondences", None)
        rows = cpu_align(scans, counts, pairs, init_transforms, **kw)
        n = len(rows)
        T = torch.zeros((n, 4, 4), dtype=torch.float64)
        T[:, 0, 3], T[:, 1, 3] = torch.from_numpy(rows[:, 0]), torch.from_numpy(rows[:, 1])
        T[:, 0, 0], T[:, 1, 0] = torch.from_numpy(np.cos(rows[:, 2])), torch.from_numpy(np.sin(rows[:, 2]))
        return engine.IcpBatchResult(T, torch.from_numpy(rows[:, 3].copy()), torch.zeros(n, dtype=torch.int32),
                                     torch.zeros(n, dtype=torch.int32), None)

    def pack_poses(T, error=None):
        out = torch.zeros((T.shape[0], 4), dtype=torch.float64)
        out[:, 0], out[:, 1] = T[:, 0, 3], T[:, 1, 3]
        out[:, 2] = torch.atan2(T[:, 1, 0], T[:, 0, 0])
        if error is not None:
            out[:, 3] = error
        return out

    engine.run_icp_batch = run_icp_batch
    engine.pack_poses = pack_poses
