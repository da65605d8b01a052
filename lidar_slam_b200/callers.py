"""Batched rewrites of the reference's three ICP caller loops (SURVEY.md section 8a row A8 /
8f row F1).  Each builds the whole pair list up front, makes ONE `run_icp_batch` launch and
finishes the cheap sequential part on the host in the reference's order.

* poses_from_scan_matching ... modules/localization.py:95-134
* loop_closures .............. main.py:158-172 (everything except the gtsam calls)
* scan_matching_with_loop_closures: both of the above in one mixed launch (config 3)
* yaw_sweep .................. code/icp_warm_up/test_icp.py:65-93
"""
from __future__ import annotations

from typing import List, Sequence

import numpy as np
import torch

from . import engine


# ---- SE(2)/SE(3) glue, same conventions as modules/utils.py:161-240 ---------------------
def T_from_poses(poses: np.ndarray) -> np.ndarray:
    """(N,3) poses (x, y, theta) -> (N,3,3) homogeneous matrices (utils.py:178-194)."""
    poses = np.asarray(poses, dtype=np.float64)
    out = np.zeros((poses.shape[0], 3, 3))
    c, s = np.cos(poses[:, 2]), np.sin(poses[:, 2])
    out[:, 0, 0], out[:, 0, 1], out[:, 0, 2] = c, -s, poses[:, 0]
    out[:, 1, 0], out[:, 1, 1], out[:, 1, 2] = s, c, poses[:, 1]
    out[:, 2, 2] = 1.0
    return out


def TSE3_from_TSE2(T2: np.ndarray) -> np.ndarray:
    """(N,3,3) -> (N,4,4) (utils.py:227-240)."""
    out = np.tile(np.eye(4), (T2.shape[0], 1, 1))
    out[:, :2, :2] = T2[:, :2, :2]
    out[:, :2, 3] = T2[:, :2, 2]
    return out


def TSE2_from_TSE3(T3: np.ndarray) -> np.ndarray:
    """(N,4,4) -> (N,3,3) (utils.py:212-225)."""
    out = np.tile(np.eye(3), (T3.shape[0], 1, 1))
    out[:, :2, :2] = T3[:, :2, :2]
    out[:, :2, 2] = T3[:, :2, 3]
    return out


def relative_inits(x_ts: np.ndarray) -> np.ndarray:
    """All T_init = inv(T(x[i-1])) @ T(x[i]), i = 1..N-1, as (N-1,4,4)
    (localization.py:116-121 with utils.py:161-176)."""
    T = T_from_poses(x_ts)
    rel = np.matmul(np.linalg.inv(T[:-1]), T[1:])
    return TSE3_from_TSE2(rel)


def _pool(z_ts: Sequence[np.ndarray]):
    return engine.pack_scans([np.asarray(z, dtype=np.float64) for z in z_ts])


def _prefix_poses(rel2: np.ndarray):
    """World poses from relative SE(2) transforms, multiplied left to right exactly like the
    reference's loop (localization.py:127-130)."""
    cur = np.eye(3)
    poses = [[0.0, 0.0, 0.0]]
    for k in range(rel2.shape[0]):
        cur = cur @ rel2[k]
        poses.append([cur[0, 2], cur[1, 2], np.arctan2(cur[1, 0], cur[0, 0])])
    return np.array(poses)


def poses_from_scan_matching(x_ts, z_ts, return_relative_poses=False):
    """Signature and results of modules/localization.py:95-134, one launch instead of N-1
    sequential run_icp calls (the pairs are independent: T_init comes from odometry)."""
    x_ts = np.asarray(x_ts, dtype=np.float64)
    n = x_ts.shape[0]
    scans, counts = _pool(z_ts)
    pairs = np.stack([np.arange(1, n), np.arange(0, n - 1)], axis=1).astype(np.int32)
    res = engine.run_icp_batch(scans, counts, pairs, relative_inits(x_ts), epsilon=0.001)
    rel2 = TSE2_from_TSE3(res.T.cpu().numpy())
    poses = _prefix_poses(rel2)
    if return_relative_poses:
        return poses, rel2
    return poses


def loop_closure_pair_list(n_poses: int, fixed_interval: int) -> np.ndarray:
    """(source i, target i+k) for i in range(0, n_poses-k, k) (main.py:161-162)."""
    i = np.arange(0, n_poses - fixed_interval, fixed_interval)
    return np.stack([i, i + fixed_interval], axis=1).astype(np.int32)


def _gate(T2: np.ndarray, max_distance: float, max_yaw_deg: float):
    angle = np.arctan2(T2[:, 1, 0], T2[:, 0, 0])
    translation = np.linalg.norm(T2[:, :2, 2], axis=1)
    return (translation < max_distance) & (np.rad2deg(angle) < max_yaw_deg)       # main.py:165-167


def loop_closures(z_ts, n_poses, fixed_interval, max_distance=np.inf, max_yaw_deg=np.inf):
    """main.py:158-172 without gtsam: returns dict(pairs, T (L,3,3), error (L,), accepted (L,) bool).
    The accepted rows are exactly the BetweenFactorPose2(i, i+k) edges the reference adds."""
    scans, counts = _pool(z_ts)
    pairs = loop_closure_pair_list(n_poses, fixed_interval)
    res = engine.run_icp_batch(scans, counts, pairs, None, epsilon=0.01, normalize_error=True)
    T2 = TSE2_from_TSE3(res.T.cpu().numpy())
    return {"pairs": pairs, "T": T2, "error": res.error.cpu().numpy(),
            "accepted": _gate(T2, max_distance, max_yaw_deg)}


def scan_matching_with_loop_closures(x_ts, z_ts, fixed_interval, max_distance=np.inf,
                                     max_yaw_deg=np.inf):
    """Config 3 (`--mode gtsam`): the N-1 odometry pairs (epsilon 1e-3, raw SSE, odometry init)
    and the loop-closure pairs (epsilon 1e-2, normalised, identity init) in ONE mixed launch,
    delivered in the reference's order (odometry factors first, then loop factors)."""
    x_ts = np.asarray(x_ts, dtype=np.float64)
    n = x_ts.shape[0]
    scans, counts = _pool(z_ts)
    odo = np.stack([np.arange(1, n), np.arange(0, n - 1)], axis=1).astype(np.int32)
    loops = loop_closure_pair_list(n, fixed_interval)
    pairs = np.concatenate([odo, loops], axis=0)
    init = np.concatenate([relative_inits(x_ts), np.tile(np.eye(4), (len(loops), 1, 1))], axis=0)
    eps = np.concatenate([np.full(len(odo), 0.001), np.full(len(loops), 0.01)])
    norm = np.concatenate([np.zeros(len(odo), dtype=bool), np.ones(len(loops), dtype=bool)])
    res = engine.run_icp_batch(scans, counts, pairs, init, epsilon=eps, normalize_error=norm)
    T2 = TSE2_from_TSE3(res.T.cpu().numpy())
    rel2, loop2 = T2[:len(odo)], T2[len(odo):]
    return {"poses": _prefix_poses(rel2), "relative_poses": rel2, "loop_pairs": loops, "loop_T": loop2,
            "loop_error": res.error.cpu().numpy()[len(odo):],
            "loop_accepted": _gate(loop2, max_distance, max_yaw_deg)}


def yaw_sweep(source_pc, target_pcs: List[np.ndarray], n_yaw=24, voxel_size=0.005, max_points=20000):
    """code/icp_warm_up/test_icp.py:65-93: per target, 24 yaw initialisations with centroid-aligned
    translation, voxel down-sampling of both clouds if either exceeds 20 000 points (done ONCE per
    target, not once per yaw), best-of-24 by error.  All targets x yaws in one launch.
    Returns (best_poses [num_pc,4,4], best_errors [num_pc] rounded to 3 dp like :93, all_errors)."""
    from .icp import voxel_downsample
    source_pc = np.asarray(source_pc, dtype=np.float64)
    clouds, pairs, inits = [], [], []
    for tgt in target_pcs:
        tgt = np.asarray(tgt, dtype=np.float64)
        src = source_pc
        if src.shape[0] > max_points or tgt.shape[0] > max_points:        # test_icp.py:75-77
            src_run, tgt_run = voxel_downsample(src, voxel_size), voxel_downsample(tgt, voxel_size)
        else:
            src_run, tgt_run = src, tgt
        si, ti = len(clouds), len(clouds) + 1
        clouds += [src_run, tgt_run]
        mu_s, mu_t = src.mean(axis=0), tgt.mean(axis=0)                    # test_icp.py:72-74 (full clouds)
        for yaw in np.linspace(0, 2 * np.pi, n_yaw, endpoint=False):
            T = np.eye(4)
            c, s = np.cos(yaw), np.sin(yaw)
            T[:3, :3] = [[c, -s, 0], [s, c, 0], [0, 0, 1]]
            T[:3, 3] = mu_t - T[:3, :3] @ mu_s
            inits.append(T)
            pairs.append([si, ti])
    scans, counts = engine.pack_scans(clouds)
    res = engine.run_icp_batch(scans, counts, np.array(pairs, dtype=np.int32), np.stack(inits),
                               epsilon=0.001, normalize_error=True)
    err = res.error.cpu().numpy().reshape(len(target_pcs), n_yaw)
    T = res.T.cpu().numpy().reshape(len(target_pcs), n_yaw, 4, 4)
    best = np.argmin(err, axis=1)                                          # first minimum, like `<` in :89
    best_poses = T[np.arange(len(target_pcs)), best]
    best_errors = np.round(err[np.arange(len(target_pcs)), best], 3)
    return best_poses, best_errors, err
