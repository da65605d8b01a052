"""Generate golden vectors for scan construction by running the UNMODIFIED reference.

Run in the authoring container only (needs /root/reference):

    python tests/golden/make_scan_golden.py

Imports /root/reference/modules/localization.py as-is (matplotlib is stubbed because
modules/utils.py imports pyplot at module level; nothing on this path plots) and stores the
inputs and the reference's outputs of get_lidar_data and statistical_outliers_removal in
tests/golden/scan_golden.npz.
"""
import contextlib
import io
import os
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

import modules.localization as ref  # noqa: E402  (the reference, unmodified)


def make_ranges(n_scans, n_beams, seed):
    """Hokuyo-shaped ranges: smooth walls + noise, 1 mm quantised through float32, with
    dropouts (0), short returns, saturated returns (> 30 m), exact gate values and NaNs."""
    rng = np.random.default_rng(seed)
    a = np.linspace(-2.356, 2.356, n_beams)
    r = 4.0 + 2.5 * np.sin(3 * a)[None, :] + rng.normal(0, 0.4, size=(n_scans, 1)) \
        + rng.normal(0, 0.01, size=(n_scans, n_beams))
    far = rng.random((n_scans, n_beams)) < 0.04
    r[far] = rng.uniform(12, 29, size=far.sum())
    r = (np.round(r * 1000) / 1000).astype(np.float32).astype(np.float64)
    r[rng.random(r.shape) < 0.01] = 0.0
    r[rng.random(r.shape) < 0.01] = 0.05
    r[rng.random(r.shape) < 0.01] = 31.5
    r[rng.random(r.shape) < 0.002] = np.nan
    r[0, :4] = [0.1, 30.0, 0.099999, 30.000001]          # the gate is inclusive on both sides
    return r


def main():
    out = {}
    for name, (n_scans, n_beams, seed) in {"a": (12, 1081, 1), "b": (5, 64, 2), "c": (3, 1080, 3)}.items():
        r = make_ranges(n_scans, n_beams, seed)
        with contextlib.redirect_stdout(io.StringIO()), contextlib.redirect_stderr(io.StringIO()):
            z = ref.get_lidar_data(r, 0.1, 30.0)
            zf = {k: ref.statistical_outliers_removal(z, k_std=k) for k in (1, 2, 3)}
        out[f"{name}/ranges"] = r
        out[f"{name}/counts"] = np.array([s.shape[0] for s in z])
        out[f"{name}/points"] = np.vstack(z)
        for k, f in zf.items():
            out[f"{name}/k{k}/counts"] = np.array([s.shape[0] for s in f])
            out[f"{name}/k{k}/points"] = np.vstack(f)
        print(name, r.shape, "valid", out[f"{name}/counts"].sum(),
              {k: int(out[f"{name}/k{k}/counts"].sum()) for k in zf})
    out["names"] = np.array(["a", "b", "c"])
    np.savez_compressed(os.path.join(HERE, "scan_golden.npz"), **out)
    print("wrote", os.path.join(HERE, "scan_golden.npz"))


if __name__ == "__main__":
    main()
