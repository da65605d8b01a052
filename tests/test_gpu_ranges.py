"""RANGES mode of the planar kernel: get_lidar_data (modules/localization.py:252-307) fused into the
staging of run_icp (modules/icp.py:123-189) -- b200icp_run_ranges / b200icp_run_host_ranges -- against
the two-step path (b200icp_scans_from_ranges, itself bit-identical to the reference's scans, then
b200icp_run), bit for bit; plus the gate's give-up path (status NOT_RESIDENT)."""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _same(a, b):
    for name in ("T", "error", "iters", "status", "corr"):
        x, y = getattr(a, name), getattr(b, name)
        if x is None:
            continue
        assert torch.equal(torch.nan_to_num(x, nan=-1.0) if x.is_floating_point() else x,
                           torch.nan_to_num(y, nan=-1.0) if y.is_floating_point() else y), name


def _workload(n_pairs, seed):
    from lidar_slam_b200 import synth
    b = synth.odometry_batch(n_pairs, seed=seed, device="cuda", with_ranges=True)
    return b["ranges"].contiguous(), b["pairs"], b["init"]


def test_ranges_entry_equals_scan_build_plus_icp_bitwise():
    import lidar_slam_b200 as eng
    from lidar_slam_b200 import engine, scans as scan_build
    ranges, pairs, init = _workload(300, 41)
    ranges[7, 100:140] = float("nan")          # NaN beams are dropped
    ranges[9, ::3] = 45.0                      # beyond the gate
    ranges[11, :] = 0.0                        # an empty scan: pairs (11, 10) and (12, 11) are not run
    pool, counts = scan_build.scans_from_ranges(ranges, 0.1, 30.0)
    eps = torch.full((300,), 1e-3, dtype=torch.float64)
    eps[::5] = 1e-2
    norm = torch.zeros(300, dtype=torch.uint8)
    norm[::5] = 1
    want = eng.run_icp_batch(pool, counts, pairs, init, epsilon=eps, normalize_error=norm,
                             return_correspondences=True)
    got, got_counts = engine.run_icp_ranges_batch(ranges, pairs, init, epsilon=eps, normalize_error=norm,
                                                  return_correspondences=True)
    _same(got, want)
    assert torch.equal(got_counts, counts)
    assert int(got.status[10]) == 4 and int(got.status[11]) == 4          # empty scan 11
    # float32 ranges: the synthetic log is millimetre-quantised through float32, so the copy is exact
    r32 = ranges.to(torch.float32)
    assert torch.equal(torch.nan_to_num(r32.double(), nan=-1.0), torch.nan_to_num(ranges, nan=-1.0))
    got32, counts32 = engine.run_icp_ranges_batch(r32, pairs, init, epsilon=eps, normalize_error=norm,
                                                  return_correspondences=True)
    _same(got32, want)
    assert torch.equal(counts32, counts)
    # the block bounds switched off: same bits
    nop, _ = engine.run_icp_ranges_batch(ranges, pairs, init, epsilon=eps, normalize_error=norm,
                                         return_correspondences=True, flags=engine.FLAG_NO_PRUNE)
    _same(nop, want)


def test_host_ranges_entry_equals_device_entry_bitwise():
    """3000 consecutive pairs from HOST ranges: one gated launch while the range rows arrive in slices."""
    import lidar_slam_b200 as eng
    from lidar_slam_b200 import _lib, engine, hostapi
    ranges, pairs, init = _workload(3000, 43)
    want, counts = engine.run_icp_ranges_batch(ranges, pairs, init, epsilon=1e-3)
    r_np, p_np, i_np = ranges.cpu().numpy(), pairs.cpu().numpy(), init.cpu().numpy()
    for flags in (0, _lib.flag_slices(1), _lib.flag_slices(63)):
        host, hcounts = hostapi.run_icp_ranges_host(r_np, p_np, i_np, epsilon=1e-3, flags=flags)
        assert np.array_equal(host.T, want.T.cpu().numpy()), flags
        assert np.array_equal(host.error, want.error.cpu().numpy())
        assert np.array_equal(host.iters, want.iters.cpu().numpy())
        assert np.array_equal(host.status, want.status.cpu().numpy())
        assert np.array_equal(hcounts, counts.cpu().numpy())
    host32, _ = hostapi.run_icp_ranges_host(r_np.astype(np.float32), p_np, i_np, epsilon=1e-3)
    assert np.array_equal(host32.T, want.T.cpu().numpy())
    perm = np.random.default_rng(1).permutation(3000)
    shuf, _ = hostapi.run_icp_ranges_host(r_np, p_np[perm], i_np[perm], epsilon=1e-3)
    assert np.array_equal(shuf.T, want.T.cpu().numpy()[perm])
    small, _ = hostapi.run_icp_ranges_host(r_np[:40], p_np[:39], i_np[:39], epsilon=1e-3)   # below the slice count
    assert np.array_equal(small.T, want.T.cpu().numpy()[:39])


def test_gated_launch_gives_up_with_status_not_resident():
    """A gate that is never raised: after the wait limit every pair is given up -- status 7, T = init,
    NaN error, workspace word 5 set -- instead of being aligned on scans that never arrived."""
    from lidar_slam_b200 import _lib, engine, synth
    lib = _lib.load()
    b = synth.odometry_batch(40, seed=5, device="cuda")
    n = 40
    eps = torch.full((n,), 1e-3, dtype=torch.float64, device="cuda")
    norm = torch.zeros(n, dtype=torch.uint8, device="cuda")
    T = torch.zeros((n, 4, 4), dtype=torch.float64, device="cuda")
    err = torch.zeros(n, dtype=torch.float64, device="cuda")
    iters = torch.full((n,), -1, dtype=torch.int32, device="cuda")
    status = torch.full((n,), -1, dtype=torch.int32, device="cuda")
    nbytes = ctypes.c_size_t()
    _lib.check(lib.b200icp_workspace_bytes(n, 1082, 2, ctypes.byref(nbytes)), "ws")
    ws = torch.zeros(nbytes.value, dtype=torch.uint8, device="cuda")
    gate = torch.tensor([20], dtype=torch.int32, device="cuda")     # scans 0..19 resident, the rest never
    p = engine._ptr
    _lib.check(lib.b200icp_run_ex(p(b["scans"]), p(b["counts"]), b["scans"].shape[0], b["scans"].shape[1], 2,
                                  p(b["pairs"]), p(b["init"]), p(eps), p(norm), n, 2000, 1e-4, p(T), p(err), p(iters),
                                  p(status), None, p(ws), ws.numel(), engine._stream_ptr(), p(gate), None, 0),
               "b200icp_run_ex")
    torch.cuda.synchronize()
    st = status.cpu().numpy()
    pairs = b["pairs"].cpu().numpy()
    resident = pairs.max(axis=1) < 20
    assert (st[resident] <= 2).all() and (st[~resident] == 7).all()
    assert torch.equal(T[~torch.from_numpy(resident).cuda()], b["init"][~torch.from_numpy(resident).cuda()])
    assert bool(torch.isnan(err[~torch.from_numpy(resident).cuda()]).all())
    assert int(ws[20:24].cpu().numpy().view(np.uint32)[0]) == 1


def test_two_submitted_calls_in_flight_equal_the_synchronous_entry():
    """b200icp_submit_host_ranges / b200icp_wait_host: two batches in flight on one device, waited for in
    either order, fill their own workspaces with the bits of the synchronous entry; a third submit is
    refused (BUSY) and a ticket cannot be waited for twice."""
    from lidar_slam_b200 import _lib, hostapi, synth
    batches = []
    for seed in (51, 52, 53):
        b = synth.odometry_batch(2300, seed=seed, device="cpu", with_ranges=True)
        r = hostapi.pinned_empty(tuple(b["ranges"].shape), np.float64); r[:] = b["ranges"].numpy()
        p = hostapi.pinned_empty((2300, 2), np.int32); p[:] = b["pairs"].numpy()
        i = hostapi.pinned_empty((2300, 4, 4), np.float64); i[:] = b["init"].numpy()
        batches.append((r, p, i))
    n_scans, n_beams = batches[0][0].shape
    want = []
    for r, p, i in batches:
        res, counts = hostapi.run_icp_ranges_host(r, p, i, epsilon=1e-3, return_correspondences=True)
        want.append((res.T.copy(), res.error.copy(), res.iters.copy(), res.correspondences.copy(), counts.copy()))
    ws = [hostapi.HostRangesWorkspace(2300, n_scans, n_beams + (n_beams & 1), True) for _ in range(2)]
    for order in ((0, 1), (1, 0)):
        for w in ws:
            w.T.fill(np.nan); w.iters.fill(-7)
        t = [hostapi.submit_icp_ranges_host(*batches[k], ws[k], epsilon=1e-3, return_correspondences=True) for k in range(2)]
        with pytest.raises(RuntimeError, match="in flight"):
            hostapi.submit_icp_ranges_host(*batches[2], ws[0], epsilon=1e-3)
        for k in order:
            res, counts = t[k].wait()
            T, err, iters, corr, cnt = want[k]
            assert np.array_equal(res.T, T) and np.array_equal(res.error, err) and np.array_equal(res.iters, iters)
            assert np.array_equal(res.correspondences, corr) and np.array_equal(counts, cnt)
    # a pipeline: submit k + 1, then wait for k
    tickets = [None, None]
    for step in range(5):
        k = step % 2
        if tickets[k] is not None:
            tickets[k].wait()
        tickets[k] = hostapi.submit_icp_ranges_host(*batches[step % 3], ws[k], epsilon=1e-3, return_correspondences=True)
    res4, _ = tickets[0].wait()           # step 4 -> batch 1
    res3, _ = tickets[1].wait()           # step 3 -> batch 0
    assert np.array_equal(res4.T, want[1][0]) and np.array_equal(res3.T, want[0][0])
    lib = _lib.load()
    assert lib.b200icp_wait_host(12345 << 8) == -1          # unknown ticket
