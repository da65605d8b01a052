"""CPU tests: the oracle against the committed golden vectors (produced by the
reference itself, tests/golden/make_golden.py) and, when mounted, the live reference."""
import numpy as np
import pytest

from conftest import golden_cases, reference_module
from oracle import icp_oracle as orc


@pytest.mark.parametrize("nn", ["kdtree", "brute"])
def test_oracle_matches_golden(golden, nn):
    for c in golden_cases(golden):
        T, err, corr, passes = orc.run_icp_info(
            c["src"], c["tgt"], c["init"], c["epsilon"], c["max_iters"],
            c["stopping_thresh"], c["normalize"], nn=nn)
        assert passes == c["passes"], c["name"]
        np.testing.assert_allclose(T, c["T"], rtol=0, atol=1e-11, err_msg=c["name"])
        np.testing.assert_allclose(err, c["err"], rtol=1e-11, err_msg=c["name"])
        assert np.array_equal(corr, c["corr"]), c["name"]


def test_return_conventions(golden):
    c = next(golden_cases(golden))
    kw = dict(init_transform=c["init"], epsilon=c["epsilon"])
    T = orc.run_icp(c["src"], c["tgt"], **kw)
    assert isinstance(T, np.ndarray) and T.shape == (4, 4)
    T2, e = orc.run_icp(c["src"], c["tgt"], return_error=True, **kw)
    T3, k = orc.run_icp(c["src"], c["tgt"], return_correspondences=True, **kw)
    T4, e4, k4 = orc.run_icp(c["src"], c["tgt"], return_error=True, return_correspondences=True, **kw)
    assert np.array_equal(T, T2) and np.array_equal(T, T3) and np.array_equal(T, T4)
    assert e == e4 and np.array_equal(k, k4) and k.dtype == np.int64 and k.shape == (c["src"].shape[0],)


def test_voxel_downsample_golden(golden):
    out = orc.voxel_downsample(golden["voxel_in"], float(golden["voxel_size"]))
    np.testing.assert_allclose(out, golden["voxel_out"], rtol=0, atol=1e-15)


def test_planar_closed_form_equals_svd_solution():
    """The closed form the device uses == LAPACK Kabsch incl. the reflection branch (icp.py:62-67)."""
    rng = np.random.default_rng(0)
    n_reflect = 0
    for _ in range(400):
        n = int(rng.integers(3, 40))
        q = np.concatenate([rng.normal(size=(n, 2)) * rng.uniform(0.1, 5, 2), np.zeros((n, 1))], axis=1)
        p = np.concatenate([rng.normal(size=(n, 2)) * rng.uniform(0.1, 5, 2), np.zeros((n, 1))], axis=1)
        full = orc.rigid_fit(q, p)
        cov = (q - q.mean(0)).T @ (p - p.mean(0))
        r2, zs = orc.planar_fit_closed_form(cov[:2, :2])
        n_reflect += zs < 0
        np.testing.assert_allclose(full[:2, :2], r2, atol=1e-12)
        assert full[2, 2] == pytest.approx(zs, abs=1e-12)
        assert np.all(full[2, :2] == 0) and np.all(full[:2, 2] == 0)
    assert n_reflect > 50


def test_brute_equals_kdtree_on_scans():
    from lidar_slam_b200 import synth
    b = synth.odometry_batch(3, seed=99)
    s, c = b["scans"].numpy(), b["counts"].numpy()
    from scipy.spatial import KDTree
    for i in range(3):
        q, t = s[i + 1, :c[i + 1]], s[i, :c[i]]
        assert np.array_equal(orc.nn_brute(q, t), KDTree(t).query(q)[1])


def test_oracle_matches_live_reference():
    ref = reference_module()
    if ref is None:
        pytest.skip("reference not mounted (expected on the GPU box)")
    from lidar_slam_b200 import synth
    b = synth.odometry_batch(6, seed=123)
    s, c, init = b["scans"].numpy(), b["counts"].numpy(), b["init"].numpy()
    for i in range(6):
        src, tgt = s[i + 1, :c[i + 1]], s[i, :c[i]]
        Tr, er, cr = ref.run_icp(src, tgt, init_transform=init[i], epsilon=1e-3,
                                 return_error=True, return_correspondences=True)
        To, eo, co = orc.run_icp(src, tgt, init_transform=init[i], epsilon=1e-3,
                                 return_error=True, return_correspondences=True, nn="brute")
        np.testing.assert_allclose(To, Tr, rtol=0, atol=1e-12)
        assert eo == pytest.approx(er, rel=1e-12) and np.array_equal(co, cr)
