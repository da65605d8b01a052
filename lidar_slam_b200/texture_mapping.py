"""Texture mapping on the device: mirror of the reference's
``modules/texture_mapping.py``.

``generate_texture_map`` keeps the reference's signature (texture_mapping.py:7-15) and result
(float32 ``(grid_map_width, grid_map_height, 3)`` in [0, 1]); the per-frame numpy pipeline is
replaced by ``b200icp_texture_paint``.  Frames are streamed to the device in chunks from pinned
host memory.  ``paint_frames`` is the array-level entry for callers that already hold the frames.
Plotting (texture_mapping.py:100-113) is out of scope.
"""
from __future__ import annotations

from typing import Callable, Optional, Tuple

import numpy as np
import torch

from . import _lib
from .engine import _device, _ptr, _stream_ptr


def closest_stamps(faster_sensor_stamps, slower_sensor_stamps) -> np.ndarray:
    """Kinect.get_closest_stamps (modules/sensors.py:138-146): for every slower stamp the index of
    the nearest faster stamp, first index on ties (np.argmin)."""
    fast = np.asarray(faster_sensor_stamps, dtype=np.float64)
    return np.array([np.abs(fast - s).argmin() for s in np.asarray(slower_sensor_stamps)], dtype=np.int64)


def world_from_robot(x_t) -> np.ndarray:
    """T_wr of texture_mapping.py:71-81 for one pose (x, y, yaw), evaluated with numpy."""
    x, y, yaw = x_t
    R_wr = np.array([
        [np.cos(yaw), -np.sin(yaw), 0],
        [np.sin(yaw), np.cos(yaw), 0],
        [0, 0, 1]
    ])
    T_wr = np.eye(4)
    T_wr[:3, :3] = R_wr
    T_wr[:3, 3] = np.array([x, y, 0])
    return T_wr


def camera_from_optical() -> np.ndarray:
    """np.linalg.inv(T_oc) of texture_mapping.py:213-222."""
    R_oc = np.array([
        [0., -1., 0.],
        [0., 0., -1.],
        [1., 0., 0.]
    ])
    T_oc = np.eye(4)
    T_oc[:3, :3] = R_oc
    return np.linalg.inv(T_oc)


class TexturePainter:
    """Device state of one texture map: a uint64 key per cell (see csrc/texture.cu)."""

    def __init__(self, width, height, res, min_x, min_y, T_rc, K, device=None):
        self.width, self.height = int(width), int(height)
        self.res, self.min_x, self.min_y = float(res), float(min_x), float(min_y)
        self.dev = torch.device(device) if device is not None else _device()
        self.keys = torch.zeros((self.width * self.height,), dtype=torch.int64, device=self.dev)
        M = np.hstack((np.asarray(K, dtype=np.float64), np.zeros((3, 1))))      # texture_mapping.py:32
        consts = np.concatenate([np.linalg.inv(M[:, :3]).ravel(),                # :191
                                 camera_from_optical().ravel(),
                                 np.asarray(T_rc, dtype=np.float64).ravel()])
        self._consts = consts                       # host copies: the C ABI reads them on the host
        self.frames_done = 0

    def paint_frames(self, disparity, rgb, poses) -> None:
        """disparity [F,h,w] float32, rgb [F,h,w,3] uint8 (RGB), poses [F,3]: frames in paint order."""
        lib = _lib.load()
        disparity = torch.as_tensor(disparity)
        rgb = torch.as_tensor(rgb)
        if disparity.dtype != torch.float32 or rgb.dtype != torch.uint8:
            raise ValueError("disparity must be float32 and rgb uint8")
        n, h, w = disparity.shape
        if rgb.shape != (n, h, w, 3):
            raise ValueError("rgb must have shape [F, h, w, 3] matching disparity")
        poses = np.asarray(poses, dtype=np.float64).reshape(n, 3)
        T_wr = np.stack([world_from_robot(p) for p in poses]) if n else np.zeros((0, 4, 4))
        d_dev = disparity.to(self.dev, non_blocking=True).contiguous()
        c_dev = rgb.to(self.dev, non_blocking=True).contiguous()
        T_dev = torch.as_tensor(np.ascontiguousarray(T_wr)).to(self.dev)
        c = self._consts
        kinv = (np.ctypeslib.as_ctypes(np.ascontiguousarray(c[:9])))
        tco = (np.ctypeslib.as_ctypes(np.ascontiguousarray(c[9:25])))
        trc = (np.ctypeslib.as_ctypes(np.ascontiguousarray(c[25:41])))
        import ctypes
        with torch.cuda.device(self.dev):
            _lib.check(lib.b200icp_texture_paint(
                _ptr(self.keys), self.width, self.height, self.res, self.min_x, self.min_y, _ptr(d_dev),
                _ptr(c_dev), n, h, w, self.frames_done, _ptr(T_dev),
                ctypes.cast(kinv, ctypes.c_void_p), ctypes.cast(tco, ctypes.c_void_p),
                ctypes.cast(trc, ctypes.c_void_p), _stream_ptr()), "b200icp_texture_paint")
        self.frames_done += n

    def resolve(self, grid_map=None) -> torch.Tensor:
        """float32 [width, height, 3] texture in [0, 1] (texture_mapping.py:46-47, :98)."""
        lib = _lib.load()
        grid = None
        if grid_map is not None:
            grid = torch.as_tensor(np.ascontiguousarray(grid_map, dtype=np.uint8)).to(self.dev) \
                if not isinstance(grid_map, torch.Tensor) else grid_map.to(self.dev, torch.uint8).contiguous()
            if grid.numel() != self.width * self.height:
                raise ValueError("grid_map must have width*height cells")
        out = torch.empty((self.width, self.height, 3), dtype=torch.float32, device=self.dev)
        with torch.cuda.device(self.dev):
            _lib.check(lib.b200icp_texture_resolve(_ptr(self.keys), _ptr(grid), _ptr(out), self.width,
                                                   self.height, _stream_ptr()), "b200icp_texture_resolve")
        return out


def read_image(path, is_disparity=False):
    """texture_mapping.py:115-130 (cv2 is imported lazily; only this loader needs it)."""
    import cv2
    if not is_disparity:
        return cv2.imread(path)[..., ::-1]
    return cv2.imread(path, cv2.IMREAD_UNCHANGED).astype(np.float32)


def _default_loader(dataset_num) -> Callable[[int, int], Tuple[np.ndarray, np.ndarray]]:
    def load(rgb_idx, disp_img_idx):
        # paths of texture_mapping.py:54-59
        disparity_path = "dataRGBD/Disparity" + str(dataset_num) + "/" + "disparity" + str(dataset_num) + \
                         "_" + str(disp_img_idx) + ".png"
        rgb_path = "dataRGBD/RGB" + str(dataset_num) + "/" + "rgb" + str(dataset_num) + "_" + \
                   str(rgb_idx + 1) + ".png"
        return read_image(disparity_path, is_disparity=True), read_image(rgb_path)
    return load


def generate_texture_map(dataset_num, x_ts, kinect, encoder, ogm, T_rc, K,
                         load_frame: Optional[Callable] = None, chunk: int = 64):
    """Same arguments and result as modules/texture_mapping.py:7-98.  `load_frame(rgb_idx,
    disp_img_idx) -> (disparity float32 [h,w], rgb uint8 [h,w,3])` defaults to the reference's
    PNG paths."""
    load = load_frame or _default_loader(dataset_num)
    pose_idx = closest_stamps(encoder.stamps, kinect.rgb_stamps)           # :35-38
    disp_idx = closest_stamps(kinect.disp_stamps, kinect.rgb_stamps)       # :41-44
    painter = TexturePainter(ogm.grid_map_width, ogm.grid_map_height, ogm.res, ogm.world_map_min_x,
                             ogm.world_map_min_y, T_rc, K)
    x_ts = np.asarray(x_ts, dtype=np.float64)
    n = int(np.asarray(kinect.rgb_stamps).shape[0])
    for lo in range(0, n, chunk):
        frames = [load(i, int(disp_idx[i])) for i in range(lo, min(lo + chunk, n))]
        disp = torch.from_numpy(np.stack([f[0] for f in frames]).astype(np.float32, copy=False)).pin_memory()
        rgb = torch.from_numpy(np.ascontiguousarray(np.stack([f[1] for f in frames]), dtype=np.uint8)).pin_memory()
        painter.paint_frames(disp, rgb, x_ts[pose_idx[lo:lo + len(frames)]])
    return painter.resolve(ogm.grid_map).cpu().numpy()
