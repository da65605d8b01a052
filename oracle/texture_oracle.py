"""CPU oracle for texture mapping -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline legs may import
this module; the product (``lidar_slam_b200``) never does.

Numpy restatement of modules/texture_mapping.py on in-memory frames:

* generate_texture_map (frame loop) ........ texture_mapping.py:7-98
* get_depth_image .......................... texture_mapping.py:132-143
* get_rgbi_rgbj ............................ texture_mapping.py:145-160
* vectorized_generate_point_cloud .......... texture_mapping.py:162-224
* transform_point_cloud / transform_points . texture_mapping.py:226-240, utils.py:140-159
* Kinect.get_closest_stamps ................ sensors.py:138-146

Parity pinning: the reference holds no golden vectors for this path; the oracle is pinned
against the UNMODIFIED generate_texture_map run on seeded synthetic frames
(tests/golden/make_texture_golden.py swaps only cv2.imread for an in-memory reader; outputs in
tests/golden/texture_golden.npz, replayed bit for bit by tests/test_texture_oracle.py).
"""
from __future__ import annotations

import numpy as np

__all__ = ["synthetic_frames", "closest_stamps", "point_cloud", "generate_texture_map", "CASES", "case_inputs",
           "T_RC", "K"]


def synthetic_frames(n_frames, h, w, seed):
    """Seeded Kinect-shaped frames: disparity float32 [F,h,w] (uint16 values, a tilted floor plus
    noise, dropouts and saturated pixels) and RGB uint8 [F,h,w,3]."""
    rng = np.random.default_rng(seed)
    rows = np.linspace(0.0, 1.0, h)[None, :, None]
    base = 640.0 + 330.0 * rows + rng.uniform(-40, 40, size=(n_frames, 1, 1))
    disp = base + 25.0 * np.sin(np.linspace(0, 6.0, w))[None, None, :] + rng.normal(0, 6.0, size=(n_frames, h, w))
    disp = np.clip(np.round(disp), 0, 2047)
    disp[rng.random(disp.shape) < 0.02] = 0
    disp[rng.random(disp.shape) < 0.01] = 1200          # dd < 0: negative depth
    rgb = rng.integers(0, 256, size=(n_frames, h, w, 3), dtype=np.uint8)
    return disp.astype(np.float32), rgb


def closest_stamps(faster, slower):
    faster = np.asarray(faster, dtype=np.float64)                          # sensors.py:24-26
    return [int(np.abs(faster - s).argmin()) for s in slower]


def transform_points(points, T):
    points = np.hstack((points, np.ones((points.shape[0], 1))))           # utils.py:153-154
    return (T @ points.T).T[:, :3]


def transform_point_cloud(pcl, T):
    out = np.zeros_like(pcl)                                               # texture_mapping.py:237-240
    out[:, :3] = transform_points(pcl[:, :3], T)
    out[:, 3:] = pcl[:, 3:]
    return out


def point_cloud(disparity_image, rgb_image, K):
    """XYZRGB cloud in the camera frame plus the flat pixel index of every kept point."""
    M = np.hstack((K, np.zeros((3, 1))))                                   # :32
    with np.errstate(divide="ignore", invalid="ignore", over="ignore"):
        dd = (-0.00304 * disparity_image) + 3.31                           # :142 (float32)
        depth_image = 1.03 / dd                                            # :143
        h, w = depth_image.shape
        i_idx, j_idx = np.meshgrid(np.arange(h), np.arange(w), indexing="ij")
        i_flat, j_flat, depth_flat = i_idx.flatten(), j_idx.flatten(), depth_image.flatten()
        K_inv = np.linalg.inv(M[:, :3])                                    # :191
        hom = np.stack([j_flat, i_flat, np.ones_like(i_flat)], axis=-1)    # :192
        xyz = (K_inv @ hom.T).T * depth_flat[:, np.newaxis]                # :193
        rgbi = ((526.37 * i_flat) + 19276 - (7877.07 * depth_flat)) / 585.051   # :158 with dd := depth (:196)
        rgbj = ((526.37 * j_flat) + 16662) / 585.051                       # :159
        valid = (0 <= rgbi) & (rgbi < h) & (0 <= rgbj) & (rgbj < w)        # :199
        vi = np.clip(rgbi[valid].astype(int), 0, h - 1)                    # :202-203
        vj = np.clip(rgbj[valid].astype(int), 0, w - 1)
    pcl = np.concatenate([xyz[valid], rgb_image[vi, vj]], axis=-1)         # :204-208
    R_oc = np.array([[0., -1., 0.], [0., 0., -1.], [1., 0., 0.]])          # :211-217
    T_oc = np.eye(4)
    T_oc[:3, :3] = R_oc
    return transform_point_cloud(pcl, np.linalg.inv(T_oc)), np.flatnonzero(valid)


def generate_texture_map(x_ts, rgb_stamps, disp_stamps, encoder_stamps, disparity, rgb, grid_map,
                         res, min_x, min_y, T_rc, K):
    """uint8 texture [W,H,3] before the final /255 and the float32 result (texture_mapping.py:46-98).
    `disparity[k]`, `rgb[k]` are the frames the reference would load for disparity index k / RGB
    index k (0-based)."""
    pose_idx = closest_stamps(encoder_stamps, rgb_stamps)                   # :35-38
    disp_idx = closest_stamps(disp_stamps, rgb_stamps)                      # :41-44
    W, H = grid_map.shape
    tex = np.repeat(np.expand_dims(grid_map, axis=2), 3, axis=2)            # :46-47
    for k in range(len(rgb_stamps)):
        x, y, yaw = x_ts[pose_idx[k]]
        pcl, _ = point_cloud(disparity[disp_idx[k]], rgb[k], K)
        pcl_r = transform_point_cloud(pcl, T_rc)                            # :68
        T_wr = np.eye(4)                                                    # :71-81
        T_wr[:3, :3] = np.array([[np.cos(yaw), -np.sin(yaw), 0], [np.sin(yaw), np.cos(yaw), 0], [0, 0, 1]])
        T_wr[:3, 3] = np.array([x, y, 0])
        pcl_w = transform_point_cloud(pcl_r, T_wr)                          # :82
        floor = pcl_w[:, [0, 1, 3, 4, 5]]                                   # :85 (no height test in the reference)
        with np.errstate(invalid="ignore"):
            gx = np.ceil((floor[:, 0] - min_x) / res).astype(np.int32) - 1  # ogm.py:118-119
            gy = np.ceil((floor[:, 1] - min_y) / res).astype(np.int32) - 1
        ok = (gx >= 0) & (gx < W) & (gy >= 0) & (gy < H)                    # :91-92
        tex[gx[ok], gy[ok], :] = floor[ok, 2:]                              # :97 last writer wins
    return tex, tex.astype(np.float32) / 255.0                              # :98


# ---- seeded cases shared by tests/golden/make_texture_golden.py and the tests ----------------
CASES = {   # name: (n_rgb, n_disp, h, w, seed, res, half_x, half_y)
    "small": (7, 9, 60, 80, 1, 0.05, 4.0, 3.0),
    "vga": (2, 3, 480, 640, 2, 0.05, 8.0, 6.0),
}


def case_inputs(name):
    n_rgb, n_disp, h, w, seed, res, hx, hy = CASES[name]
    rng = np.random.default_rng(100 + seed)
    disparity, _ = synthetic_frames(n_disp + 1, h, w, seed)            # index 0 unused (paths are what the stamps say)
    _, rgb = synthetic_frames(n_rgb, h, w, seed + 50)
    rgb_stamps = np.sort(rng.uniform(0, 10, n_rgb))
    disp_stamps = np.sort(rng.uniform(0, 10, n_disp + 1))
    enc_stamps = np.linspace(0, 10, 40)
    t = enc_stamps / 10
    x_ts = np.stack([1.5 * np.cos(2 * t), 1.0 * np.sin(2 * t), 2 * t + 0.3], axis=1)
    W = int(np.ceil(2 * hx / res + 1))
    H = int(np.ceil(2 * hy / res + 1))
    grid = (rng.random((W, H)) < 0.5).astype(np.uint8)
    return dict(disparity=disparity, rgb=rgb, rgb_stamps=rgb_stamps, disp_stamps=disp_stamps,
                enc_stamps=enc_stamps, x_ts=x_ts, grid=grid, res=res, min_x=-hx, min_y=-hy)


T_RC = np.eye(4)                                                             # main.py:217-225
T_RC[:3, :3] = np.array([[np.cos(np.deg2rad(18.)), 0, np.sin(np.deg2rad(18.))], [0, 1, 0],
                         [-np.sin(np.deg2rad(18.)), 0, np.cos(np.deg2rad(18.))]])
T_RC[:3, 3] = np.array([0.16766, 0., 0.38001])
K = np.array([[585.05, 0, 242.94], [0, 585.05, 315.84], [0, 0, 1]])          # main.py:228-232
