"""The drop-in under the reference's OWN caller: lidar_slam_b200.icp.install() makes
`from .icp import run_icp` in the reference's modules/localization.py resolve to the B200 engine,
then the reference's unmodified poses_from_scan_matching (modules/localization.py:95-134) runs its
sequential loop over it.  Compared with the batched rewrite (callers.poses_from_scan_matching) and
with the oracle.  Skipped unless the reference tree is reachable ($LIDAR_SLAM_REF, /root/reference,
baseline/_ref) -- it never is copied into the repository."""
import os
import sys
import types

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _reference_root():
    for root in (os.environ.get("LIDAR_SLAM_REF"), "/root/reference", os.path.join(ROOT, "baseline", "_ref")):
        if root and os.path.isfile(os.path.join(root, "modules", "localization.py")):
            return root
    return None


@pytest.mark.skipif(_reference_root() is None, reason="reference tree not reachable on this box")
def test_reference_poses_from_scan_matching_runs_on_the_dropin():
    import torch
    from lidar_slam_b200 import callers, synth
    import lidar_slam_b200.icp as dropin
    ref_root = _reference_root()
    saved = {k: sys.modules.get(k) for k in ("matplotlib", "matplotlib.pyplot", "modules", "modules.icp",
                                             "modules.localization", "modules.utils", "icp")}
    try:
        for name in ("matplotlib", "matplotlib.pyplot"):          # plotting only; not installed here
            if name not in sys.modules:
                sys.modules[name] = types.ModuleType(name)
        sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
        for k in ("modules", "modules.localization", "modules.utils"):
            sys.modules.pop(k, None)
        sys.path.insert(0, ref_root)
        dropin.install()                                           # BEFORE modules.localization is imported
        import modules.localization as ref_loc                     # the unmodified reference module
        assert ref_loc.run_icp is dropin.run_icp
        b = synth.odometry_batch(50, seed=7, device="cpu")
        scans, counts = b["scans"].numpy(), b["counts"].numpy()
        z_ts = [scans[i, :counts[i]] for i in range(51)]
        poses = b["poses"][0]
        rng = np.random.default_rng(3)
        x_ts = poses + rng.normal(scale=[0.01, 0.01, 0.003], size=poses.shape)      # odometry-like input poses
        got_poses, got_rel = ref_loc.poses_from_scan_matching(x_ts, z_ts, True)
        want_poses, want_rel = callers.poses_from_scan_matching(x_ts, z_ts, True)
        # same kernels, but a batch of one uses the latency CTA shape (16 warps): another summation order
        assert np.abs(np.asarray(got_rel) - want_rel).max() < 1e-11
        assert np.abs(np.asarray(got_poses) - want_poses).max() < 1e-10
        # and against the oracle driving the same loop
        from oracle import icp_oracle as orc
        rel = callers.relative_inits(x_ts)
        for i in (0, 17, 49):
            To, _, _, _ = orc.run_icp_info(z_ts[i + 1], z_ts[i], rel[i], epsilon=0.001)
            assert np.abs(callers.TSE2_from_TSE3(To[None])[0] - np.asarray(got_rel)[i]).max() < 1e-9
    finally:
        if ref_root in sys.path:
            sys.path.remove(ref_root)
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
