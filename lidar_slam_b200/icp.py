"""Drop-in replacement for the reference's `modules/icp.py` (and its duplicate
`code/icp_warm_up/icp.py`): same public names, keyword names, defaults and return
conventions, computed by the B200 engine.

    import lidar_slam_b200.icp as icp          # or install(), see INTEGRATION.md
    T = icp.run_icp(z[i], z[i-1], init_transform=T0, epsilon=0.001)

Every call is a batch of one through the same kernels `run_icp_batch` uses
(`b200icp_run_pair_host`: one pinned staging block, one copy up, one launch with a
latency-oriented CTA shape, one copy back; no torch on the path), and is synchronous
like the reference.
"""
from __future__ import annotations

import sys

import numpy as np

from . import hostapi

__all__ = ["run_icp", "voxel_downsample", "install"]


def _cloud(pc, name):
    if hasattr(pc, "detach"):                 # torch tensors are accepted as an extension
        pc = pc.detach().cpu().numpy()
    pc = np.ascontiguousarray(pc, dtype=np.float64)
    if pc.ndim != 2 or pc.shape[1] not in (2, 3):
        raise ValueError(f"{name} must have shape (N,2) or (N,3), got {pc.shape}")
    if pc.shape[0] == 0:
        raise ValueError(f"{name} is empty")   # the reference fails inside scipy/numpy here
    return pc


def run_icp(
    pc1,
    pc2,
    init_transform=np.eye(4),
    epsilon=0.01,
    max_iters=2000,
    stopping_thresh=0.0001,
    return_error=False,
    normalize_error=False,
    return_correspondences=False,
):
    """Signature and semantics of modules/icp.py:123-189.

    Returns T (4,4) float64 alone, or (T, error), (T, correspondences),
    (T, error, correspondences) by the flags; `error` and `correspondences` are
    those of the last executed pass; correspondences are int64 of shape (N_s,).
    """
    src, tgt = _cloud(pc1, "pc1"), _cloud(pc2, "pc2")
    if src.shape[1] != tgt.shape[1]:      # the reference pads 2-D clouds with z = 0 (icp.py:153-156)
        pad = lambda c: c if c.shape[1] == 3 else np.ascontiguousarray(np.concatenate([c, np.zeros((c.shape[0], 1))], axis=1))
        src, tgt = pad(src), pad(tgt)
    init = np.ascontiguousarray(init_transform, dtype=np.float64)
    if init.shape != (4, 4):
        raise ValueError("init_transform must be 4x4")
    T, error, _, _, corr32 = hostapi.run_icp_pair_host(src, tgt, init, float(epsilon), int(max_iters),
                                                       float(stopping_thresh), bool(normalize_error),
                                                       bool(return_correspondences))
    error = np.float64(error)
    if return_correspondences:
        corr = corr32.astype(np.int64)
    if return_error and not return_correspondences:
        return T, error
    if return_correspondences and not return_error:
        return T, corr
    if return_error and return_correspondences:
        return T, error, corr
    return T


def voxel_downsample(point_cloud, voxel_size):
    """Signature and semantics of modules/icp.py:4-27 (per-voxel mean, voxels in
    lexicographic index order), computed on the device."""
    import torch
    from . import engine
    pts = _cloud(point_cloud, "point_cloud")
    if pts.shape[1] != 3:
        pts3 = np.concatenate([pts, np.zeros((pts.shape[0], 1))], axis=1)
        out = engine.voxel_downsample_device(torch.from_numpy(pts3).to(engine._device()), voxel_size)
        return out[:, :2].cpu().numpy()
    out = engine.voxel_downsample_device(torch.from_numpy(pts).to(engine._device()), voxel_size)
    return out.cpu().numpy()


def install(names=("modules.icp", "icp")):
    """Make `from modules.icp import run_icp` / `from icp import run_icp` resolve to this
    module.  Call BEFORE importing modules.localization, which binds run_icp at import
    time (modules/localization.py:5); see INTEGRATION.md."""
    me = sys.modules[__name__]
    for n in names:
        sys.modules[n] = me
    return me
