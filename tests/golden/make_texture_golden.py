"""Generate golden texture maps by running the UNMODIFIED reference.

Run in the authoring container only (needs /root/reference and cv2):

    python tests/golden/make_texture_golden.py

Imports /root/reference/modules/texture_mapping.py as-is (matplotlib stubbed), replaces ONLY
`cv2.imread` with an in-memory reader that serves the seeded synthetic frames of
oracle.texture_oracle.synthetic_frames (the dataRGBD/ PNGs are not shipped), runs the
reference's generate_texture_map with its own Kinect class and stores the result.  Inputs are
regenerated from their seeds by the tests; only small inputs and the outputs are stored.
"""
import contextlib
import io
import os
import re
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.environ.get("LIDAR_SLAM_REF", "/root/reference"))
for name in ("matplotlib", "matplotlib.pyplot"):
    sys.modules.setdefault(name, types.ModuleType(name))
sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]

import modules.texture_mapping as ref  # noqa: E402  (the reference, unmodified)
from modules.sensors import Kinect  # noqa: E402
from oracle import texture_oracle as to  # noqa: E402

CASES, case_inputs, T_RC, K = to.CASES, to.case_inputs, to.T_RC, to.K


def main():
    out = {}
    for name in CASES:
        c = case_inputs(name)

        def imread(path, flag=None, c=c):
            k = int(re.search(r"_(\d+)\.png$", path).group(1))
            if "disparity" in path:
                return c["disparity"][k].astype(np.uint16)                   # the reference casts to float32
            return np.ascontiguousarray(c["rgb"][k - 1][..., ::-1])          # BGR on disk; rgb index is 1-based

        ref.cv2.imread = imread
        kinect = Kinect({"disp_stamps": c["disp_stamps"], "rgb_stamps": c["rgb_stamps"]})
        enc = types.SimpleNamespace(stamps=c["enc_stamps"])
        W, H = c["grid"].shape
        ogm = types.SimpleNamespace(grid_map=c["grid"], grid_map_width=W, grid_map_height=H)
        res, mnx, mny = c["res"], c["min_x"], c["min_y"]

        def world2grid(x, y):                                                # modules/ogm.py:118-120, verbatim arithmetic
            x_g = np.ceil((x - mnx) / res).astype(np.int32) - 1
            y_g = np.ceil((y - mny) / res).astype(np.int32) - 1
            return np.hstack((x_g.reshape(-1, 1), y_g.reshape(-1, 1)))
        ogm.world2grid = world2grid
        with contextlib.redirect_stdout(io.StringIO()), contextlib.redirect_stderr(io.StringIO()), \
                np.errstate(all="ignore"):
            tex = ref.generate_texture_map(20, c["x_ts"], kinect, enc, ogm, T_RC, K)
        u8 = np.round(tex * 255).astype(np.uint8)
        assert tex.dtype == np.float32 and np.array_equal(u8.astype(np.float32) / 255.0, tex)
        out[f"{name}/texture_u8"] = u8
        painted = (u8 != np.repeat(c["grid"][..., None], 3, axis=2)).any(axis=2).sum()
        print(name, tex.shape, "cells that changed", int(painted))
    np.savez_compressed(os.path.join(HERE, "texture_golden.npz"), **out)
    print("wrote", os.path.join(HERE, "texture_golden.npz"))


if __name__ == "__main__":
    main()
