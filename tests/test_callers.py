"""Batched caller rewrites (lidar_slam_b200/callers.py) against the reference's loops restated
around the oracle: localization.py:112-130, main.py:161-172, test_icp.py:65-93."""
import numpy as np
import pytest

from oracle import icp_oracle as orc


def _T_from_pose(p):            # modules/utils.py:178-194
    x, y, th = p
    return np.array([[np.cos(th), -np.sin(th), x], [np.sin(th), np.cos(th), y], [0, 0, 1]])


def _se3(T2):                   # modules/utils.py:227-240
    T = np.eye(4)
    T[:2, :2] = T2[:2, :2]
    T[:2, 3] = T2[:2, 2]
    return T


def _se2(T3):                   # modules/utils.py:212-225
    T = np.eye(3)
    T[:2, :2] = T3[:2, :2]
    T[:2, 2] = T3[:2, 3]
    return T


def _problem(n, seed):
    from lidar_slam_b200 import synth
    b = synth.odometry_batch(n, seed=seed)
    s, c = b["scans"].numpy(), b["counts"].numpy()
    z_ts = [s[i, :c[i]] for i in range(n + 1)]
    rng = np.random.default_rng(seed)
    x_ts = b["poses"][0] + rng.normal(scale=[0.004, 0.004, 0.001], size=(n + 1, 3))   # "odometry"
    return x_ts, z_ts


def test_relative_inits_match_reference_formula():
    from lidar_slam_b200 import callers
    rng = np.random.default_rng(0)
    x = np.cumsum(rng.normal(scale=[0.02, 0.02, 0.01], size=(50, 3)), axis=0)
    got = callers.relative_inits(x)
    for i in range(1, 50):
        want = _se3(np.dot(np.linalg.inv(_T_from_pose(x[i - 1])), _T_from_pose(x[i])))   # utils.py:161-176
        np.testing.assert_allclose(got[i - 1], want, rtol=0, atol=1e-15)
    assert np.array_equal(callers.loop_closure_pair_list(47, 10),
                          np.array([[i, i + 10] for i in range(0, 37, 10)]))


@pytest.mark.gpu
def test_poses_from_scan_matching_equals_reference_loop():
    from lidar_slam_b200 import callers
    x_ts, z_ts = _problem(24, seed=41)
    poses, rel = callers.poses_from_scan_matching(x_ts, z_ts, return_relative_poses=True)
    cur = _T_from_pose([0, 0, 0])
    want_poses, want_rel = [[0, 0, 0]], []
    for i in range(1, len(x_ts)):                                   # localization.py:115-130
        T_init = np.dot(np.linalg.inv(_T_from_pose(x_ts[i - 1])), _T_from_pose(x_ts[i]))
        T = _se2(orc.run_icp(z_ts[i], z_ts[i - 1], init_transform=_se3(T_init), epsilon=0.001))
        want_rel.append(T)
        cur = cur @ T
        want_poses.append([cur[0, 2], cur[1, 2], np.arctan2(cur[1, 0], cur[0, 0])])
    np.testing.assert_allclose(rel, np.array(want_rel), rtol=0, atol=1e-9)
    np.testing.assert_allclose(poses, np.array(want_poses), rtol=0, atol=1e-8)
    assert poses.shape == (25, 3) and rel.shape == (24, 3, 3)


@pytest.mark.gpu
def test_mixed_launch_equals_two_reference_loops():
    from lidar_slam_b200 import callers
    x_ts, z_ts = _problem(31, seed=43)
    out = callers.scan_matching_with_loop_closures(x_ts, z_ts, fixed_interval=10, max_distance=0.5,
                                                   max_yaw_deg=5.0)
    lc = callers.loop_closures(z_ts, len(x_ts), 10, max_distance=0.5, max_yaw_deg=5.0)
    assert np.array_equal(out["loop_pairs"], lc["pairs"])
    np.testing.assert_array_equal(out["loop_T"], lc["T"])
    np.testing.assert_array_equal(out["loop_error"], lc["error"])
    for k, i in enumerate(range(0, len(x_ts) - 10, 10)):            # main.py:161-167
        T, err = orc.run_icp(z_ts[i], z_ts[i + 10], return_error=True, normalize_error=True)
        T2 = _se2(T)
        np.testing.assert_allclose(out["loop_T"][k], T2, rtol=0, atol=1e-9)
        assert out["loop_error"][k] == pytest.approx(err, rel=1e-9)
        ok = np.linalg.norm(T2[:2, 2]) < 0.5 and np.rad2deg(np.arctan2(T2[1, 0], T2[0, 0])) < 5.0
        assert bool(out["loop_accepted"][k]) == bool(ok)
    poses = callers.poses_from_scan_matching(x_ts, z_ts)
    np.testing.assert_array_equal(out["poses"], poses)


@pytest.mark.gpu
def test_yaw_sweep_with_voxel_downsampling():
    from lidar_slam_b200 import callers, synth
    src = synth.container_cloud(21000, seed=3)                       # > 20000 -> both clouds down-sampled
    tgts = [synth.partial_view(src, seed=4), synth.partial_view(src, seed=5)]
    best_T, best_e, all_e = callers.yaw_sweep(src, tgts, voxel_size=0.01)
    assert best_T.shape == (2, 4, 4) and all_e.shape == (2, 24)
    for j, tgt in enumerate(tgts):
        s_ds, t_ds = orc.voxel_downsample(src, 0.01), orc.voxel_downsample(tgt, 0.01)
        for k in (0, 7, 16):                                         # spot-check three of the 24 inits
            yaw = np.linspace(0, 2 * np.pi, 24, endpoint=False)[k]
            T0 = np.eye(4)
            T0[:3, :3] = [[np.cos(yaw), -np.sin(yaw), 0], [np.sin(yaw), np.cos(yaw), 0], [0, 0, 1]]
            T0[:3, 3] = tgt.mean(0) - T0[:3, :3] @ src.mean(0)
            _, e = orc.run_icp(s_ds, t_ds, init_transform=T0, epsilon=0.001, return_error=True,
                               normalize_error=True)
            assert all_e[j, k] == pytest.approx(e, rel=1e-8)
        assert best_e[j] == round(float(all_e[j].min()), 3)
