"""CPU: the texture-mapping oracle replays, bit for bit, the maps the UNMODIFIED reference
produced on the seeded synthetic frames (tests/golden/texture_golden.npz)."""
import os

import numpy as np
import pytest

from oracle import texture_oracle as to

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def texture_golden():
    return np.load(os.path.join(ROOT, "tests", "golden", "texture_golden.npz"), allow_pickle=False)


@pytest.mark.parametrize("name", list(to.CASES))
def test_oracle_replays_reference_texture(texture_golden, name):
    c = to.case_inputs(name)
    u8, tex = to.generate_texture_map(c["x_ts"], c["rgb_stamps"], c["disp_stamps"], c["enc_stamps"],
                                      c["disparity"], c["rgb"], c["grid"], c["res"], c["min_x"], c["min_y"],
                                      to.T_RC, to.K)
    assert tex.dtype == np.float32
    assert np.array_equal(u8, texture_golden[f"{name}/texture_u8"])
    assert np.array_equal(tex, texture_golden[f"{name}/texture_u8"].astype(np.float32) / 255.0)


def test_closest_stamps_first_index_on_ties():
    from lidar_slam_b200 import texture_mapping as tm
    fast = np.array([0.0, 1.0, 2.0, 3.0])
    slow = np.array([0.5, 2.5, -1.0, 9.0, 1.0])
    assert to.closest_stamps(fast, slow) == [0, 2, 0, 3, 1]
    assert tm.closest_stamps(fast, slow).tolist() == [0, 2, 0, 3, 1]
