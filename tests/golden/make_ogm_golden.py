"""Generate golden occupancy-grid vectors by running the UNMODIFIED reference.

Run in the authoring container only (needs /root/reference):

    python tests/golden/make_ogm_golden.py

Imports /root/reference/modules/ogm.py as-is (matplotlib stubbed: the plotting helpers are not
called) and stores seeded inputs (poses, ragged scans) with the reference's float32 log-odds
grid and uint8 occupancy grid in tests/golden/ogm_golden.npz.
"""
import contextlib
import io
import os
import sys
import types

import numpy as np

# The reference updates a float32 grid with `grid[idx] -= np.float64(...)`: under NumPy >= 2 (NEP 50)
# that is f32(f64(cell) - log 4), which is what csrc/ogm.cu reproduces; NumPy 1.x value-based casting would
# evaluate it in float32.  The fixtures are therefore tied to the NumPy major version that made them.
assert int(np.__version__.split('.')[0]) >= 2, 'regenerate the ogm goldens under NumPy >= 2 (NEP 50 promotion)'

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.environ.get("LIDAR_SLAM_REF", "/root/reference"))
for name in ("matplotlib", "matplotlib.pyplot"):
    sys.modules.setdefault(name, types.ModuleType(name))
sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]

from modules.ogm import OccupancyGridMap  # noqa: E402  (the reference, unmodified)
from lidar_slam_b200 import synth  # noqa: E402


def case_room(n_scans, every, seed):
    """Scans of the synthetic room along a trajectory, thinned to every `every`-th beam."""
    poses = synth.make_trajectories(1, n_scans, seed)[0]
    poses[:, :2] *= 0.6                                      # stay well inside the room
    scans, counts = synth.cast_scans(poses, seed)
    scans, counts = scans.numpy(), counts.numpy()
    return poses, [scans[i, :counts[i]][::every] for i in range(n_scans)]


def main():
    out = {}
    cases = {
        # name: (res, (max_x, max_y, min_x, min_y), poses, scans)
        "room": (0.05, (9.0, 6.0, -9.0, -6.0)) + case_room(12, 5, 3),
        # small map: many rays leave the grid (clipping of the in-bounds part, ogm.py:175-183)
        "clipped": (0.1, (3.0, 2.0, -2.5, -2.0)) + case_room(8, 7, 4),
        # coarse cells: many rays share cells, end cells are crossed by other rays (mixed order)
        "coarse": (0.5, (9.0, 6.0, -9.0, -6.0)) + case_room(40, 3, 5),
    }
    # saturation: the same scan 30 times drives cells to the +/-20 clip (ogm.py:191)
    p, z = case_room(1, 4, 6)
    cases["saturate"] = (0.25, (9.0, 6.0, -9.0, -6.0), np.repeat(p, 30, axis=0), z * 30)
    for name, (res, (mxx, mxy, mnx, mny), poses, scans) in cases.items():
        ogm = OccupancyGridMap(res, mxx, mxy, mnx, mny)
        with contextlib.redirect_stdout(io.StringIO()), contextlib.redirect_stderr(io.StringIO()):
            ogm.build_map(poses, scans)
        out[f"{name}/params"] = np.array([res, mxx, mxy, mnx, mny])
        out[f"{name}/poses"] = poses
        out[f"{name}/counts"] = np.array([s.shape[0] for s in scans])
        out[f"{name}/points"] = np.vstack(scans)
        out[f"{name}/log_odds"] = ogm.grid_map_log_odds
        out[f"{name}/grid"] = ogm.grid_map
        lo = ogm.grid_map_log_odds
        print(name, lo.shape, "touched", int((lo != 0).sum()), "occupied", int(ogm.grid_map.sum()),
              "min/max", lo.min(), lo.max())
    out["names"] = np.array(list(cases))
    np.savez_compressed(os.path.join(HERE, "ogm_golden.npz"), **out)
    print("wrote", os.path.join(HERE, "ogm_golden.npz"))


if __name__ == "__main__":
    main()
