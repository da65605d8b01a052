"""Synthetic inputs shaped like the reference's data (which is not shipped).

The reference's datasets (`data/Hokuyo20.npz`, `data/<obj>/model.mat`) are
git-ignored upstream, so every configuration runs on synthetic inputs
(SURVEY.md section 0.1 / 8d):

* 2-D: a Hokuyo UTM-30LX style scanner (1081 beams over -135..+135 deg,
  modules/localization.py:273-278) ray-cast in a fixed polygonal room, with
  1 cm range noise, 1 mm quantisation through float32 (as
  code/test_ranges.npy), ~0.5 % invalid beams, the [0.1, 30] m gate
  (localization.py:288) and the lidar->robot offset 0.13323 m
  (localization.py:297).  Motions follow the statistics of
  outputs/relative_poses_odom_20.npy.
* 3-D: surface samples of a "drill"-like solid (cylinder + box) and partial
  views of it, for the 24-yaw sweep of code/icp_warm_up/test_icp.py:70-91.

The 2-D generator is written with torch ops so that the 10^6-pair
configuration can be generated on the GPU a rank owns; on CPU it gives the
same numbers for the same seed (float64 throughout, CPU generator for the
random draws so results do not depend on the device).
"""
from __future__ import annotations

import math

import numpy as np
import torch

N_BEAMS = 1081
ANGLE_MIN = -135.0 * math.pi / 180.0
ANGLE_MAX = 135.0 * math.pi / 180.0
LIDAR_FORWARD = 0.13323          # localization.py:297
RANGE_MIN, RANGE_MAX = 0.1, 30.0  # Hokuyo gate used by main.py:84

# World: outer room 14 x 9 m and three interior boxes, as closed polygons.
_ROOM = [(-7.0, -4.5), (7.0, -4.5), (7.0, 4.5), (-7.0, 4.5)]
_BOXES = [
    [(-1.0, -0.6), (1.2, -0.6), (1.2, 0.7), (-1.0, 0.7)],        # centre block
    [(-6.2, 2.9), (-5.0, 2.9), (-5.0, 3.9), (-6.2, 3.9)],        # corner cabinet
    [(5.1, -4.0), (6.3, -4.0), (6.3, -3.2), (5.1, -3.2)],        # corner crate
]


def world_segments():
    """(M,4) float64 array of wall segments x0,y0,x1,y1."""
    segs = []
    for poly in [_ROOM] + _BOXES:
        for a, b in zip(poly, poly[1:] + poly[:1]):
            segs.append((a[0], a[1], b[0], b[1]))
    return np.asarray(segs, dtype=np.float64)


def make_trajectories(n_traj, n_steps, seed):
    """Robot poses (n_traj, n_steps, 3) walking an ellipse around the centre block.

    Arc-length step ~ U(0, 0.035) m, heading = path tangent + N(0, 0.004) rad
    jitter, which reproduces the per-step motion statistics of the reference's
    odometry (|t| <= 0.035 m, |dtheta| ~ 5e-3 rad).
    """
    rng = np.random.default_rng(seed)
    ax = rng.uniform(4.2, 5.2, size=(n_traj, 1))
    ay = rng.uniform(2.3, 3.0, size=(n_traj, 1))
    phase0 = rng.uniform(0.0, 2.0 * np.pi, size=(n_traj, 1))
    step = rng.uniform(0.0, 0.035, size=(n_traj, n_steps))
    step[:, 0] = 0.0
    # d(arc) ~= mean radius * d(phase)
    phase = phase0 + np.cumsum(step / (0.5 * (ax + ay)), axis=1)
    x = ax * np.cos(phase)
    y = ay * np.sin(phase)
    heading = np.arctan2(ay * np.cos(phase), -ax * np.sin(phase))
    heading = heading + np.clip(rng.normal(0.0, 0.004, size=heading.shape), -0.02, 0.02)
    return np.stack([x, y, heading], axis=-1)


def cast_scans(poses, seed, device="cpu", chunk=2048, noise_sigma=0.01,
               invalid_frac=0.005, return_ranges=False):
    """Ray-cast one scan per pose.

    poses: (S,3) array-like of robot poses (x, y, theta) in the world.
    Returns (scans[S,1081,2] float64, counts[S] int32) on `device`; the valid
    points of each scan are packed to the front in beam order, the tail is 0.
    return_ranges: also return the measured ranges [S,1081] float64 (what the
    reference's Hokuyo log holds: millimetre-quantised, float32-representable,
    dropped beams 0), the input of get_lidar_data (localization.py:252-307).
    """
    poses = torch.as_tensor(np.asarray(poses), dtype=torch.float64)
    n_scans = poses.shape[0]
    gen = torch.Generator(device="cpu").manual_seed(int(seed))
    seg = torch.as_tensor(world_segments(), device=device)
    sx, sy = seg[:, 0], seg[:, 1]
    ex, ey = seg[:, 2] - seg[:, 0], seg[:, 3] - seg[:, 1]
    beam = torch.linspace(ANGLE_MIN, ANGLE_MAX, N_BEAMS, dtype=torch.float64, device=device)
    scans = torch.zeros((n_scans, N_BEAMS, 2), dtype=torch.float64, device=device)
    counts = torch.zeros((n_scans,), dtype=torch.int32, device=device)
    ranges = torch.zeros((n_scans, N_BEAMS), dtype=torch.float64, device=device) if return_ranges else None
    for lo in range(0, n_scans, chunk):
        p = poses[lo:lo + chunk].to(device)
        m = p.shape[0]
        noise = torch.randn((m, N_BEAMS), generator=gen, dtype=torch.float64).to(device)
        drop = (torch.rand((m, N_BEAMS), generator=gen, dtype=torch.float64) < invalid_frac).to(device)
        ox = p[:, 0] + LIDAR_FORWARD * torch.cos(p[:, 2])
        oy = p[:, 1] + LIDAR_FORWARD * torch.sin(p[:, 2])
        ang = p[:, 2:3] + beam[None, :]
        dx, dy = torch.cos(ang), torch.sin(ang)                       # (m, beams)
        # ray o + r d  vs segment s + u e :  r = cross(s-o, e)/cross(d, e)
        wx = sx[None, None, :] - ox[:, None, None]
        wy = sy[None, None, :] - oy[:, None, None]
        den = dx[..., None] * ey - dy[..., None] * ex
        r = (wx * ey - wy * ex) / den
        u = (wx * dy[..., None] - wy * dx[..., None]) / den
        hit = (r > 0) & (u >= 0) & (u <= 1) & (den != 0)
        r = torch.where(hit, r, torch.full_like(r, float("inf")))
        rng_true = r.min(dim=-1).values
        meas = rng_true + noise_sigma * noise
        meas = (torch.round(meas * 1000.0) / 1000.0).to(torch.float32).to(torch.float64)
        meas = torch.where(drop, torch.zeros_like(meas), meas)
        if ranges is not None:
            ranges[lo:lo + m] = meas
        valid = (meas >= RANGE_MIN) & (meas <= RANGE_MAX)
        px = meas * torch.cos(beam)[None, :] + LIDAR_FORWARD
        py = meas * torch.sin(beam)[None, :]
        order = torch.argsort((~valid).to(torch.int8), dim=1, stable=True)
        px = torch.gather(px, 1, order)
        py = torch.gather(py, 1, order)
        n_valid = valid.sum(dim=1)
        keep = torch.arange(N_BEAMS, device=device)[None, :] < n_valid[:, None]
        scans[lo:lo + m, :, 0] = torch.where(keep, px, torch.zeros_like(px))
        scans[lo:lo + m, :, 1] = torch.where(keep, py, torch.zeros_like(py))
        counts[lo:lo + m] = n_valid.to(torch.int32)
    if return_ranges:
        return scans, counts, ranges
    return scans, counts


def relative_motions(poses):
    """(S,3) poses -> (S-1,4,4) SE(3)-embedded inv(T_{i-1}) T_i.

    This is the `T_init` of the reference's scan-matching loop
    (localization.py:116-121 with utils.py:161-176,227-240).
    """
    poses = np.asarray(poses, dtype=np.float64)
    a, b = poses[:-1], poses[1:]
    dth = b[:, 2] - a[:, 2]
    ca, sa = np.cos(a[:, 2]), np.sin(a[:, 2])
    dx, dy = b[:, 0] - a[:, 0], b[:, 1] - a[:, 1]
    out = np.tile(np.eye(4), (poses.shape[0] - 1, 1, 1))
    out[:, 0, 0] = np.cos(dth)
    out[:, 0, 1] = -np.sin(dth)
    out[:, 1, 0] = np.sin(dth)
    out[:, 1, 1] = np.cos(dth)
    out[:, 0, 3] = ca * dx + sa * dy
    out[:, 1, 3] = -sa * dx + ca * dy
    return out


def perturb_planar(transforms, seed, sigma_t=0.01, sigma_r=0.003):
    """Odometry-like error on top of true motions."""
    rng = np.random.default_rng(seed)
    out = np.array(transforms, dtype=np.float64, copy=True)
    n = out.shape[0]
    th = np.arctan2(out[:, 1, 0], out[:, 0, 0]) + rng.normal(0.0, sigma_r, n)
    out[:, 0, 0], out[:, 0, 1] = np.cos(th), -np.sin(th)
    out[:, 1, 0], out[:, 1, 1] = np.sin(th), np.cos(th)
    out[:, 0, 3] += rng.normal(0.0, sigma_t, n)
    out[:, 1, 3] += rng.normal(0.0, sigma_t, n)
    return out


def odometry_batch(n_pairs, seed=20, device="cpu", n_traj=None, with_ranges=False):
    """Config-2 / config-5 shaped workload.

    Returns dict(scans[S,1081,2], counts[S], pairs[B,2] (src,tgt), init[B,4,4]).
    Pair i aligns scan i+1 (source) to scan i (target) inside one trajectory,
    initialised with perturbed true motion, like localization.py:112-123.
    with_ranges: also "ranges" [S,1081] float64, the lidar log the scans come from.
    """
    if n_traj is None:
        n_traj = max(1, n_pairs // 4955)
    per = -(-n_pairs // n_traj)            # pairs per trajectory
    poses = make_trajectories(n_traj, per + 1, seed)
    cast = cast_scans(poses.reshape(-1, 3), seed + 1, device=device, return_ranges=with_ranges)
    scans, counts = cast[0], cast[1]
    src = (np.arange(n_traj)[:, None] * (per + 1) + np.arange(1, per + 1)[None, :]).reshape(-1)
    pairs = np.stack([src, src - 1], axis=1)[:n_pairs].astype(np.int32)
    init = np.concatenate([relative_motions(poses[t]) for t in range(n_traj)], axis=0)[:n_pairs]
    init = perturb_planar(init, seed + 2)
    out = {
        "scans": scans, "counts": counts,
        "pairs": torch.as_tensor(pairs, device=device),
        "init": torch.as_tensor(init, device=device),
        "poses": poses,
    }
    if with_ranges:
        out["ranges"] = cast[2]
    return out


def thin_scans(scans, counts, seed, lo=100, hi=N_BEAMS):
    """Ragged, shortened scans standing in for the DBSCAN/range filters
    (localization.py:201-250), which leave n_i in roughly [100, 1081]."""
    rng = np.random.default_rng(seed)
    scans = scans.clone()
    counts = counts.clone()
    for i in range(scans.shape[0]):
        n = int(counts[i])
        keep_n = int(rng.integers(min(lo, n), min(hi, n) + 1))
        idx = np.sort(rng.choice(n, size=keep_n, replace=False))
        kept = scans[i, torch.as_tensor(idx, device=scans.device)]
        scans[i].zero_()
        scans[i, :keep_n] = kept
        counts[i] = keep_n
    return scans, counts


def thin_scans_fast(scans, counts, seed, lo=100):
    """Vectorised thin_scans (its own random stream): every scan keeps a uniformly drawn number of
    beams in [min(lo, n), n], chosen uniformly, beam order kept.  numpy in, numpy out."""
    scans = np.asarray(scans)
    counts = np.asarray(counts).astype(np.int64)
    S, N = scans.shape[0], scans.shape[1]
    rng = np.random.default_rng(seed)
    lo_i = np.minimum(lo, counts)
    keep_n = lo_i + np.floor(rng.random(S) * (counts - lo_i + 1)).astype(np.int64)
    keep_n = np.minimum(keep_n, counts)
    keys = rng.random((S, N))
    keys[np.arange(N)[None, :] >= counts[:, None]] = 2.0            # not a beam
    rank = np.argsort(np.argsort(keys, axis=1, kind="stable"), axis=1, kind="stable")
    keep = rank < keep_n[:, None]
    order = np.argsort(~keep, axis=1, kind="stable")               # kept beams first, beam order preserved
    out = np.take_along_axis(scans, order[:, :, None], axis=1)
    out[np.arange(N)[None, :] >= keep_n[:, None]] = 0.0
    return out, keep_n.astype(np.int32)


def gtsam_batch(n_pairs=4955, fixed_interval=10, seed=20, device="cpu"):
    """Config-3 shaped workload (`main.py --mode gtsam --filter_lidar --fixed_interval 10`): ragged
    filtered scans; the consecutive-scan pairs (odometry init, epsilon 1e-3, raw SSE;
    localization.py:112-123) followed by the loop-closure pairs (i, i + k), i = 0, k, 2k, ... with
    identity init, epsilon 1e-2 and normalised error (main.py:161-172), as ONE pair list.
    Returns dict(scans, counts, pairs, init, epsilon, normalize, n_odometry)."""
    b = odometry_batch(n_pairs, seed=seed, device=device, n_traj=1)       # ray casting where the caller computes
    scans, counts = thin_scans_fast(b["scans"].cpu().numpy(), b["counts"].cpu().numpy(), seed + 5)
    loop_src = np.arange(0, n_pairs + 1 - fixed_interval, fixed_interval)
    pairs = np.concatenate([b["pairs"].cpu().numpy(), np.stack([loop_src, loop_src + fixed_interval], axis=1).astype(np.int32)])
    init = np.concatenate([b["init"].cpu().numpy(), np.tile(np.eye(4), (len(loop_src), 1, 1))])
    eps = np.concatenate([np.full(n_pairs, 1e-3), np.full(len(loop_src), 1e-2)])
    norm = np.concatenate([np.zeros(n_pairs, np.uint8), np.ones(len(loop_src), np.uint8)])
    to = lambda x: torch.as_tensor(x).to(device)
    return {"scans": to(scans), "counts": to(counts), "pairs": to(pairs), "init": to(init), "epsilon": to(eps),
            "normalize": to(norm), "n_odometry": n_pairs}


def yaw_sweep_batch(n_source, n_targets=4, n_yaw=24, seed=1, kind="drill"):
    """Configs 1 / 4 (code/icp_warm_up/test_icp.py:61-91): one canonical model, n_targets partial views,
    n_yaw yaw initialisations per view with centroid-aligned translation.  numpy arrays:
    dict(clouds [source, targets...], pairs [n_targets*n_yaw, 2], init [.., 4, 4])."""
    src = drill_cloud(n_source, seed=seed) if kind == "drill" else container_cloud(n_source, seed=seed)
    tgts = [partial_view(src, seed=10 * seed + k) for k in range(n_targets)]
    pairs, inits = [], []
    for j, t in enumerate(tgts):
        for T0 in yaw_sweep_inits(src, t, n_yaw):
            pairs.append((0, 1 + j))
            inits.append(T0)
    return {"clouds": [src] + tgts, "pairs": np.array(pairs, np.int32), "init": np.stack(inits)}


# --------------------------------------------------------------------------
# 3-D objects
# --------------------------------------------------------------------------
def drill_cloud(n_points, seed):
    """Surface samples of a cylinder (body) + box (battery) + thin cylinder (bit)."""
    rng = np.random.default_rng(seed)
    n_body = int(0.45 * n_points)
    n_box = int(0.35 * n_points)
    n_bit = n_points - n_body - n_box
    th = rng.uniform(0, 2 * np.pi, n_body)
    body = np.stack([0.025 * np.cos(th), rng.uniform(-0.09, 0.06, n_body), 0.025 * np.sin(th) + 0.10], axis=1)
    face = rng.integers(0, 3, n_box)
    uvw = rng.uniform(-1, 1, (n_box, 3))
    sign = np.where(rng.random(n_box) < 0.5, -1.0, 1.0)
    uvw[np.arange(n_box), face] = sign
    box = uvw * np.array([0.035, 0.045, 0.04]) + np.array([0.0, -0.05, 0.0])
    th = rng.uniform(0, 2 * np.pi, n_bit)
    bit = np.stack([0.006 * np.cos(th), rng.uniform(0.06, 0.14, n_bit), 0.006 * np.sin(th) + 0.10], axis=1)
    return np.concatenate([body, box, bit], axis=0)


def container_cloud(n_points, seed):
    """Capped cylinder ("liq_container")."""
    rng = np.random.default_rng(seed)
    n_side = int(0.7 * n_points)
    n_cap = n_points - n_side
    th = rng.uniform(0, 2 * np.pi, n_side)
    side = np.stack([0.05 * np.cos(th), 0.05 * np.sin(th), rng.uniform(0, 0.2, n_side)], axis=1)
    r = 0.05 * np.sqrt(rng.uniform(0, 1, n_cap))
    th = rng.uniform(0, 2 * np.pi, n_cap)
    cap = np.stack([r * np.cos(th), r * np.sin(th), np.where(rng.random(n_cap) < 0.5, 0.0, 0.2)], axis=1)
    return np.concatenate([side, cap], axis=0)


def partial_view(cloud, seed, keep=0.55, noise=0.0005):
    """A camera-like partial view: half-space visibility, random yaw + shift, noise."""
    rng = np.random.default_rng(seed)
    view = rng.normal(size=3)
    view /= np.linalg.norm(view)
    depth = (cloud - cloud.mean(axis=0)) @ view
    vis = cloud[depth <= np.quantile(depth, keep)]
    yaw = rng.uniform(0, 2 * np.pi)
    rot = np.array([[np.cos(yaw), -np.sin(yaw), 0.0], [np.sin(yaw), np.cos(yaw), 0.0], [0.0, 0.0, 1.0]])
    shift = rng.uniform(-0.3, 0.3, 3) + np.array([0.0, 0.0, 0.6])
    out = vis @ rot.T + shift + rng.normal(0.0, noise, vis.shape)
    return out[rng.permutation(out.shape[0])]


def yaw_sweep_inits(source, target, n_yaw=24):
    """(n_yaw,4,4) initial transforms of test_icp.py:70-74: yaw about z, then
    translation aligning the centroids."""
    out = np.tile(np.eye(4), (n_yaw, 1, 1))
    mu_s, mu_t = source.mean(axis=0), target.mean(axis=0)
    for k, yaw in enumerate(np.linspace(0, 2 * np.pi, n_yaw, endpoint=False)):
        c, s = np.cos(yaw), np.sin(yaw)
        out[k, :3, :3] = [[c, -s, 0], [s, c, 0], [0, 0, 1]]
        out[k, :3, 3] = mu_t - out[k, :3, :3] @ mu_s
    return out
