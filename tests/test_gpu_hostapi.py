"""GPU parity of the host-buffer entry b200icp_run_host (numpy in, numpy out, no torch on the
path): equal to the device-pointer entry bit for bit, equal to the oracle, pipelined and
single-chunk, shuffled pair lists, non-planar lifting, optional outputs."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
TIGHT = 1e-9


def _pose_delta(Ta, Tb):
    return float(np.abs(Ta[:3, :3] - Tb[:3, :3]).max()), float(np.abs(Ta[:3, 3] - Tb[:3, 3]).max())


def _batch(n, seed):
    from lidar_slam_b200 import synth
    b = synth.odometry_batch(n, seed=seed)
    return {k: (v.numpy() if hasattr(v, "numpy") else v) for k, v in b.items()}


def test_host_entry_equals_oracle_small_batch():
    from lidar_slam_b200 import hostapi
    from oracle import icp_oracle as orc
    b = _batch(48, 7)
    res = hostapi.run_icp_batch_host(b["scans"], b["counts"], b["pairs"], b["init"], epsilon=1e-3,
                                     return_correspondences=True)
    for k in range(48):
        s, t = b["pairs"][k]
        src, tgt = b["scans"][s, :b["counts"][s]], b["scans"][t, :b["counts"][t]]
        To, eo, co, po = orc.run_icp_info(src, tgt, b["init"][k], epsilon=1e-3)
        assert res.iters[k] == po
        rot, trans = _pose_delta(res.T[k], To)
        assert rot < TIGHT and trans < TIGHT
        assert res.error[k] == pytest.approx(eo, rel=1e-9)
        assert np.array_equal(res.correspondences[k, :b["counts"][s]], co)


def test_pipelined_host_entry_equals_device_entry_bitwise():
    """3000 consecutive pairs: ONE gated launch while the pool arrives in slices (the kernel waits
    per pair on the device-side counter of resident scans); the same pairs in shuffled order (the
    first pairs then wait for most of the pool); and the chunked path (FLAG_HOST_CHUNKED: six
    launches over two streams) that the general kernel still uses."""
    import torch
    import lidar_slam_b200 as eng
    from lidar_slam_b200 import hostapi
    b = _batch(3000, 12)
    dev = eng.run_icp_batch(torch.from_numpy(b["scans"]).cuda(), torch.from_numpy(b["counts"]).cuda(),
                            torch.from_numpy(b["pairs"]).cuda(), torch.from_numpy(b["init"]).cuda(), epsilon=1e-3)
    host = hostapi.run_icp_batch_host(b["scans"], b["counts"], b["pairs"], b["init"], epsilon=1e-3)
    assert np.array_equal(host.T, dev.T.cpu().numpy())
    assert np.array_equal(host.error, dev.error.cpu().numpy())
    assert np.array_equal(host.iters, dev.iters.cpu().numpy())
    assert np.array_equal(host.status, dev.status.cpu().numpy())
    perm = np.random.default_rng(0).permutation(3000)
    shuf = hostapi.run_icp_batch_host(b["scans"], b["counts"], b["pairs"][perm], b["init"][perm], epsilon=1e-3)
    assert np.array_equal(shuf.T, host.T[perm]) and np.array_equal(shuf.iters, host.iters[perm])
    again = hostapi.run_icp_batch_host(b["scans"], b["counts"], b["pairs"], b["init"], epsilon=1e-3)
    assert np.array_equal(again.T, host.T)          # arena reuse
    from lidar_slam_b200 import _lib
    chunked = hostapi.run_icp_batch_host(b["scans"], b["counts"], b["pairs"], b["init"], epsilon=1e-3,
                                         flags=_lib.FLAG_HOST_CHUNKED)
    assert np.array_equal(chunked.T, host.T) and np.array_equal(chunked.iters, host.iters)
    for slices in (1, 3, 63, 200):
        r = hostapi.run_icp_batch_host(b["scans"], b["counts"], b["pairs"], b["init"], epsilon=1e-3,
                                       flags=_lib.flag_slices(slices))
        assert np.array_equal(r.T, host.T) and np.array_equal(r.error, host.error), slices


def test_host_entry_lifts_nonplanar_inits_and_handles_3d():
    from lidar_slam_b200 import hostapi, synth
    from oracle import icp_oracle as orc
    b = _batch(4, 4)
    init = b["init"].copy()
    init[1, 2, 3] = 0.05                                  # one pair leaves the plane: the batch is lifted
    res = hostapi.run_icp_batch_host(b["scans"], b["counts"], b["pairs"], init, epsilon=1e-3)
    for k in range(4):
        s, t = b["pairs"][k]
        To, eo, _, po = orc.run_icp_info(b["scans"][s, :b["counts"][s]], b["scans"][t, :b["counts"][t]],
                                         init[k], epsilon=1e-3)
        rot, trans = _pose_delta(res.T[k], To)
        assert rot < TIGHT and trans < TIGHT and res.iters[k] == po
        assert res.error[k] == pytest.approx(eo, rel=1e-9)
    src = synth.drill_cloud(700, seed=3)
    tgt = synth.partial_view(src, seed=5)
    pool, counts = hostapi.pack_scans_numpy([src, tgt])
    inits = np.stack(synth.yaw_sweep_inits(src, tgt, 24)[:3])
    r3 = hostapi.run_icp_batch_host(pool, counts, np.array([[0, 1]] * 3), inits, epsilon=1e-3, normalize_error=True)
    for k in range(3):
        To, eo, _, po = orc.run_icp_info(src, tgt, inits[k], epsilon=1e-3, normalize_error=True)
        rot, trans = _pose_delta(r3.T[k], To)
        assert rot < 1e-8 and trans < 1e-8 and r3.iters[k] == po


def test_host_entry_statuses_and_argument_errors():
    from lidar_slam_b200 import hostapi
    b = _batch(3, 2)
    counts = b["counts"].copy()
    counts[0] = 0                                          # empty target for pair 0 (pairs are (i, i-1))
    pairs = np.array([[1, 0], [2, 1], [-1, 2], [2, 99]], dtype=np.int32)
    init = np.broadcast_to(np.eye(4), (4, 4, 4)).copy()
    res = hostapi.run_icp_batch_host(b["scans"], counts, pairs, init, epsilon=1e-3)
    assert list(res.status[[0, 2, 3]]) == [4, 6, 6] and res.status[1] in (0, 2)
    assert res.iters[0] == 0 and np.isnan(res.error[0]) and np.array_equal(res.T[0], np.eye(4))
    with pytest.raises(ValueError):
        hostapi.run_icp_batch_host(np.zeros((2, 3, 3)), counts[:2], pairs[:1], init[:1])   # odd row length
    with pytest.raises(ValueError):
        hostapi.run_icp_batch_host(b["scans"], counts, pairs, init[:2])
    hostapi.release()


def test_plain_c_caller_recovers_a_known_motion(tmp_path):
    """tests/c/abi_smoke.c: a C99 program (no Python, no CUDA headers) aligns two scans through
    b200icp_run_host and checks the transform it gets back."""
    import subprocess
    from test_host import _build_c_caller
    exe = _build_c_caller(tmp_path)
    out = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout + out.stderr


def test_host_entry_from_two_threads():
    """ctypes drops the GIL during the call; the library serialises calls on one device (shared
    arena and streams).  Two threads, different batches, results equal the single-threaded ones."""
    import threading
    from lidar_slam_b200 import hostapi
    a, b = _batch(600, 21), _batch(2500, 22)
    want_a = hostapi.run_icp_batch_host(a["scans"], a["counts"], a["pairs"], a["init"], epsilon=1e-3)
    want_b = hostapi.run_icp_batch_host(b["scans"], b["counts"], b["pairs"], b["init"], epsilon=1e-3)
    got = {}

    def work(name, d, reps):
        for _ in range(reps):
            got[name] = hostapi.run_icp_batch_host(d["scans"], d["counts"], d["pairs"], d["init"], epsilon=1e-3)

    ts = [threading.Thread(target=work, args=("a", a, 6)), threading.Thread(target=work, args=("b", b, 3))]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    assert np.array_equal(got["a"].T, want_a.T) and np.array_equal(got["a"].iters, want_a.iters)
    assert np.array_equal(got["b"].T, want_b.T) and np.array_equal(got["b"].iters, want_b.iters)


def test_integration_md_numpy_binding_snippet_runs():
    """The ctypes binding printed in INTEGRATION.md section 3a is executed as it stands (only the
    library path is made absolute) and must reproduce the package's own result."""
    import os
    import re
    from lidar_slam_b200 import _lib, hostapi
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    text = open(os.path.join(root, "INTEGRATION.md")).read()
    blocks = re.findall(r"```python\n(.*?)```", text, flags=re.S)
    snippet = [b for b in blocks if "b200icp_run_host" in b and "def run_icp_pairs" in b]
    assert len(snippet) == 1
    code = snippet[0].replace('"libb200icp.so"', repr(_lib.LIB_PATH))
    ns = {}
    exec(compile(code, "INTEGRATION.md:3a", "exec"), ns)
    b = _batch(40, 5)
    z_ts = [b["scans"][i, :b["counts"][i]] for i in range(b["scans"].shape[0])]
    pairs = [tuple(p) for p in b["pairs"]]
    T, err = ns["run_icp_pairs"](z_ts, pairs, b["init"], epsilon=1e-3)
    want = hostapi.run_icp_batch_host(b["scans"], b["counts"], b["pairs"], b["init"], epsilon=1e-3)
    assert np.array_equal(T, want.T) and np.array_equal(err, want.error)
