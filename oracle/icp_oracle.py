"""CPU oracle for the ICP hot path -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` /
``--impl reference`` legs of ``bench.py`` may import this module.  The product
(``lidar_slam_b200``) never does; it fails loudly when its CUDA library is
missing.

This is a float64 numpy/scipy restatement, written from the algorithm
description in SURVEY.md section 3.4, of the reference's point-to-point ICP:

* driver loop, stop-test order, return tuple ........ modules/icp.py:123-189
* one pass (transform, match, solve, compose, SSE) ... modules/icp.py:100-121
* nearest neighbour (scipy KD-tree, k=1, p=2) ........ modules/icp.py:29-41,161
* rigid solve (Kabsch via 3x3 SVD + det fix) ......... modules/icp.py:43-74
* SSE / bbox-normalised error ........................ modules/icp.py:76-98
* voxel down-sampling ................................ modules/icp.py:4-27

Parity pinning: the reference holds NO golden vectors for this path
(SURVEY.md section 8c), so the oracle is pinned against outputs of the
reference itself: ``tests/golden/make_golden.py`` imports
``/root/reference/modules/icp.py`` in the authoring container and commits its
inputs/outputs under ``tests/golden/``; ``tests/test_oracle.py`` replays them
through this file (and, when ``/root/reference`` is present, compares live).

Two matchers are offered: ``nn="kdtree"`` uses the same third-party routine as
the reference (scipy.spatial.KDTree; scipy is unpinned upstream, 1.18.1 here)
and is what the CPU baseline times; ``nn="brute"`` is an exhaustive float64
search (lowest index wins ties) and is what the CUDA engine is specified to
equal.  They agree except at exact distance ties.
"""
from __future__ import annotations

import numpy as np
from scipy.spatial import KDTree

__all__ = [
    "run_icp", "run_icp_info", "voxel_downsample", "nn_brute", "rigid_fit",
    "alignment_error", "planar_fit_closed_form",
]


# --------------------------------------------------------------------------
# building blocks
# --------------------------------------------------------------------------
def _as_xyz(cloud):
    """(N,2)|(N,3) -> (N,3) float64, planar input gets z = 0 (icp.py:153-156)."""
    cloud = np.asarray(cloud, dtype=np.float64)
    if cloud.ndim != 2 or cloud.shape[1] not in (2, 3):
        raise ValueError("point cloud must have shape (N,2) or (N,3)")
    if cloud.shape[1] == 2:
        cloud = np.concatenate([cloud, np.zeros((cloud.shape[0], 1))], axis=1)
    return cloud


def nn_brute(queries, targets, chunk=256):
    """Exact float64 1-NN by exhaustive search; ties -> lowest target index.

    Specification of what the CUDA matcher must return (icp.py:40 semantics,
    tie rule ours).  Distances are formed as sum over axes of squared
    differences, like the KD-tree's leaf scan.
    """
    queries = np.asarray(queries, np.float64)
    targets = np.asarray(targets, np.float64)
    out = np.empty(queries.shape[0], dtype=np.int64)
    for lo in range(0, queries.shape[0], chunk):
        q = queries[lo:lo + chunk]
        d = np.zeros((q.shape[0], targets.shape[0]))
        for ax in range(q.shape[1]):
            diff = q[:, ax, None] - targets[None, :, ax]
            d += diff * diff
        out[lo:lo + chunk] = np.argmin(d, axis=1)
    return out


def rigid_fit(moving, fixed):
    """Least-squares rotation+translation taking ``moving`` onto ``fixed``.

    Kabsch with the determinant correction placed on the third singular
    direction, exactly as icp.py:55-72 does it (so planar data with a
    negative-determinant 2x2 covariance yields a reflection in the plane and
    -1 on the z axis).  Returns a 4x4.
    """
    mu_m = moving.mean(axis=0)
    mu_f = fixed.mean(axis=0)
    cov = (moving - mu_m).T @ (fixed - mu_f)
    u, _, vt = np.linalg.svd(cov)
    v = vt.T
    fix = np.eye(3)
    fix[2, 2] = np.linalg.det(v @ u.T)
    rot = v @ fix @ u.T
    out = np.eye(4)
    out[:3, :3] = rot
    out[:3, 3] = mu_f - rot @ mu_m
    return out


def planar_fit_closed_form(cov2):
    """Closed-form top-left 2x2 of ``rigid_fit`` for planar data.

    ``cov2`` is the 2x2 block of the cross-covariance sum_j (q_j-qbar)(p_j-pbar)^T.
    det >= 0: proper rotation by atan2(c01-c10, c00+c11); det < 0: in-plane
    reflection [[c,s],[s,-c]] with angle atan2(c01+c10, c00-c11) and -1 on z.
    This mirrors what the device code computes; tests check it against
    ``rigid_fit`` (i.e. against LAPACK).  Returns (2x2 matrix, z sign).
    """
    a, b, c, d = cov2[0, 0], cov2[0, 1], cov2[1, 0], cov2[1, 1]
    if a * d - b * c >= 0.0:
        num, den = b - c, a + d
        h = np.hypot(num, den)
        co, si = (1.0, 0.0) if h == 0.0 else (den / h, num / h)
        return np.array([[co, -si], [si, co]]), 1.0
    num, den = b + c, a - d
    h = np.hypot(num, den)
    co, si = den / h, num / h
    return np.array([[co, si], [si, -co]]), -1.0


def alignment_error(moved, matched, normalize=False):
    """Sum of squared residuals; optionally / (bbox_diag^2 * N) (icp.py:88-97)."""
    sse = np.sum((moved - matched) ** 2)
    if not normalize:
        return sse
    both = np.concatenate([moved, matched], axis=0)
    extent = both.max(axis=0) - both.min(axis=0)
    diag = np.sqrt(np.sum(extent ** 2))
    return sse / (diag ** 2 * moved.shape[0])


# --------------------------------------------------------------------------
# the driver
# --------------------------------------------------------------------------
def run_icp_info(pc1, pc2, init_transform=None, epsilon=0.01, max_iters=2000,
                 stopping_thresh=0.0001, normalize_error=False, nn="kdtree"):
    """Full-information ICP: returns (T, error, correspondences, passes).

    ``passes`` is the number of executed passes (>= 1).  ``error`` and
    ``correspondences`` belong to the LAST executed pass, i.e. they describe
    the transform *before* the final update (icp.py:119-120,183-189).
    """
    src = _as_xyz(pc1)
    tgt = _as_xyz(pc2)
    if src.shape[0] == 0 or tgt.shape[0] == 0:
        raise ValueError("empty point cloud")
    pose = np.eye(4) if init_transform is None else np.array(init_transform, dtype=np.float64)
    tree = KDTree(tgt) if nn == "kdtree" else None
    src_h = np.concatenate([src, np.ones((src.shape[0], 1))], axis=1)

    prev_err = None
    count = 0          # the reference's `iteration`
    passes = 0
    while True:
        moved = (pose @ src_h.T).T[:, :3]
        if tree is not None:
            match = tree.query(moved)[1]
        else:
            match = nn_brute(moved, tgt)
        picked = tgt[match]
        step = rigid_fit(moved, picked)
        err = alignment_error(moved, picked, normalize=normalize_error)
        pose = step @ pose
        passes += 1
        # stop tests, in the reference's order (icp.py:172-181)
        if err < epsilon:
            break
        if count >= max_iters:
            break
        if prev_err is not None and abs(prev_err - err) < stopping_thresh:
            break
        prev_err = err
        count += 1
    return pose, err, np.asarray(match, dtype=np.int64), passes


def run_icp(pc1, pc2, init_transform=None, epsilon=0.01, max_iters=2000,
            stopping_thresh=0.0001, return_error=False, normalize_error=False,
            return_correspondences=False, nn="kdtree"):
    """Same signature and return convention as the reference (icp.py:123-189)."""
    pose, err, match, _ = run_icp_info(
        pc1, pc2, init_transform, epsilon, max_iters, stopping_thresh,
        normalize_error, nn=nn)
    if return_error and return_correspondences:
        return pose, err, match
    if return_error:
        return pose, err
    if return_correspondences:
        return pose, match
    return pose


def voxel_downsample(point_cloud, voxel_size):
    """Per-voxel mean, voxels in lexicographic index order (icp.py:4-27)."""
    pts = np.asarray(point_cloud, dtype=np.float64)
    keys = np.floor((pts - pts.min(axis=0)) / voxel_size).astype(int)
    uniq, inverse = np.unique(keys, axis=0, return_inverse=True)
    inverse = np.asarray(inverse).reshape(-1)
    out = np.empty((uniq.shape[0], pts.shape[1]))
    for v in range(uniq.shape[0]):
        out[v] = np.mean(pts[inverse == v], axis=0)
    return out
