"""GPU parity of texture mapping through the C ABI: bit-identical to the
reference's maps (golden) and to the oracle; last-writer-wins order; streaming in chunks."""
import os
import types

import numpy as np
import pytest
import torch

from oracle import texture_oracle as to

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def texture_golden():
    return np.load(os.path.join(ROOT, "tests", "golden", "texture_golden.npz"), allow_pickle=False)


def _run_drop_in(c, chunk):
    from lidar_slam_b200 import texture_mapping as tm
    W, H = c["grid"].shape
    kinect = types.SimpleNamespace(rgb_stamps=c["rgb_stamps"], disp_stamps=c["disp_stamps"])
    enc = types.SimpleNamespace(stamps=c["enc_stamps"])
    ogm = types.SimpleNamespace(grid_map=c["grid"], grid_map_width=W, grid_map_height=H, res=c["res"],
                                world_map_min_x=c["min_x"], world_map_min_y=c["min_y"])
    load = lambda rgb_idx, disp_idx: (c["disparity"][disp_idx], c["rgb"][rgb_idx])
    return tm.generate_texture_map(20, c["x_ts"], kinect, enc, ogm, to.T_RC, to.K, load_frame=load, chunk=chunk)


@pytest.mark.parametrize("name,chunk", [("small", 64), ("small", 2), ("vga", 64)])
def test_texture_bitwise_vs_reference_golden(texture_golden, name, chunk):
    c = to.case_inputs(name)
    tex = _run_drop_in(c, chunk)
    want = texture_golden[f"{name}/texture_u8"].astype(np.float32) / 255.0
    assert tex.dtype == np.float32 and tex.shape == want.shape
    bad = np.argwhere((tex != want).any(axis=2))
    assert bad.shape[0] == 0, (name, bad[:5])


def test_last_writer_wins_and_unpainted_cells_keep_the_grid():
    """Same frame painted twice with different colours: the second wins everywhere; cells no
    pixel reaches keep the grid value on all channels."""
    from lidar_slam_b200 import texture_mapping as tm
    c = to.case_inputs("small")
    W, H = c["grid"].shape
    p = tm.TexturePainter(W, H, c["res"], c["min_x"], c["min_y"], to.T_RC, to.K)
    d = c["disparity"][1:2]
    rgb_a = np.full((1,) + d.shape[1:] + (3,), 10, dtype=np.uint8)
    rgb_b = np.full_like(rgb_a, 200)
    pose = c["x_ts"][:1]
    p.paint_frames(d, rgb_a, pose)
    p.paint_frames(d, rgb_b, pose)
    tex = p.resolve(c["grid"]).cpu().numpy()
    painted = (p.keys != 0).cpu().numpy().reshape(W, H)
    assert painted.sum() > 50
    assert np.array_equal(tex[painted], np.full((painted.sum(), 3), np.float32(200) / np.float32(255)))
    g = c["grid"].astype(np.float32) / 255.0
    assert np.array_equal(tex[~painted], np.repeat(g[~painted][:, None], 3, axis=1))


def test_full_size_oracle_spot_check():
    """Fresh 480x640 frames (not in the golden file): device == oracle."""
    from lidar_slam_b200 import texture_mapping as tm
    disparity, rgb = to.synthetic_frames(3, 480, 640, seed=77)
    stamps = np.array([0.0, 1.0, 2.0])
    x_ts = np.array([[0.2, -0.1, 0.4], [0.5, 0.3, -1.2], [-0.7, 0.6, 2.9]])
    grid = np.zeros((241, 201), dtype=np.uint8)
    u8, want = to.generate_texture_map(x_ts, stamps, stamps, stamps, disparity, rgb, grid, 0.05, -6.0, -5.0,
                                       to.T_RC, to.K)
    p = tm.TexturePainter(241, 201, 0.05, -6.0, -5.0, to.T_RC, to.K)
    p.paint_frames(disparity, rgb, x_ts)
    tex = p.resolve(grid).cpu().numpy()
    assert np.array_equal(tex, want)
