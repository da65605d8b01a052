"""b200icp_run_static: the allocation-free variant of b200icp_run (scratch in a fixed device array
inside the library; include/b200icp.h).  It has no counterpart in the reference
-- so the check is against the existing entry point: the same
kernels run, hence the outputs are compared bit for bit (stricter than the 1e-9 the parity tests
use), on consecutive calls that leave different stale contents in the shared array.

Collected last on purpose (file name)."""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else None


def _problem_3d(n_src, n_targets, n_yaw, seed):
    import lidar_slam_b200 as eng
    from lidar_slam_b200 import synth
    src = synth.drill_cloud(n_src, seed=seed)
    tgts = [synth.partial_view(src, seed=10 * seed + k) for k in range(n_targets)]
    pool, counts = eng.pack_scans([src] + tgts)
    pairs, inits = [], []
    for j, t in enumerate(tgts):
        for T0 in synth.yaw_sweep_inits(src, t, n_yaw):
            pairs.append((0, 1 + j))
            inits.append(T0)
    return pool, counts, np.array(pairs, np.int32), np.stack(inits), 3


def _problem_2d(n_scans, seed):
    from lidar_slam_b200 import synth
    b = synth.odometry_batch(n_scans, seed=seed)
    return b["scans"], b["counts"], b["pairs"].numpy().astype(np.int32), b["init"].numpy(), 2


class _Device:
    """Inputs and one set of outputs on the GPU (allocated BEFORE the calls under test)."""

    def __init__(self, pool, counts, pairs, inits, dim, epsilon, normalize):
        dev = torch.device("cuda:0")
        self.dim = dim
        self.pool = torch.as_tensor(pool).to(dev, torch.float64).contiguous()
        self.counts = torch.as_tensor(counts).to(dev, torch.int32).contiguous()
        self.pairs = torch.as_tensor(pairs).to(dev, torch.int32).contiguous()
        self.inits = torch.as_tensor(inits).to(dev, torch.float64).contiguous()
        self.B = int(pairs.shape[0])
        self.n_scans, self.n_max = int(self.pool.shape[0]), int(self.pool.shape[1])
        self.eps = torch.full((self.B,), epsilon, dtype=torch.float64, device=dev)
        self.norm = torch.full((self.B,), 1 if normalize else 0, dtype=torch.uint8, device=dev)

    def outputs(self):
        dev = self.pool.device
        return dict(T=torch.full((self.B, 4, 4), float("nan"), dtype=torch.float64, device=dev),
                    err=torch.full((self.B,), float("nan"), dtype=torch.float64, device=dev),
                    iters=torch.full((self.B,), -7, dtype=torch.int32, device=dev),
                    status=torch.full((self.B,), -7, dtype=torch.int32, device=dev),
                    corr=torch.full((self.B, self.n_max), -7, dtype=torch.int32, device=dev))

    def head(self):
        return [_ptr(self.pool), _ptr(self.counts), self.n_scans, self.n_max, self.dim, _ptr(self.pairs),
                _ptr(self.inits), _ptr(self.eps), _ptr(self.norm), self.B, 60, 1e-9]


def _run_existing(lib, d, out):
    need = ctypes.c_size_t(0)
    assert lib.b200icp_workspace_bytes(d.B, d.n_max, d.dim, ctypes.byref(need)) == 0
    ws = torch.empty(max(need.value, 256), dtype=torch.uint8, device=d.pool.device)
    rc = lib.b200icp_run(*d.head(), _ptr(out["T"]), _ptr(out["err"]), _ptr(out["iters"]), _ptr(out["status"]),
                         _ptr(out["corr"]), _ptr(ws), ws.numel(), None)
    assert rc == 0, lib.b200icp_error_string(rc)
    torch.cuda.synchronize()
    return need.value


def _run_static(lib, d, out):
    rc = lib.b200icp_run_static(*d.head(), _ptr(out["T"]), _ptr(out["err"]), _ptr(out["iters"]),
                                _ptr(out["status"]), _ptr(out["corr"]), None)
    assert rc == 0, lib.b200icp_error_string(rc)
    torch.cuda.synchronize()


def _assert_same(a, b):
    for k in ("T", "err", "iters", "status", "corr"):
        x, y = a[k], b[k]
        if x.is_floating_point():
            assert torch.equal(torch.isnan(x), torch.isnan(y)), k
            x, y = torch.nan_to_num(x, nan=0.0), torch.nan_to_num(y, nan=0.0)
        assert torch.equal(x, y), k
    assert int((a["iters"] >= 1).sum()) == a["iters"].numel()       # every pair really ran


def test_static_scratch_entry_matches_b200icp_run_on_consecutive_calls():
    from lidar_slam_b200 import _lib
    lib = _lib.load()
    problems = [
        _Device(*_problem_3d(1500, 2, 6, seed=2), epsilon=1e-3, normalize=True),      # general kernel
        _Device(*_problem_3d(2600, 3, 4, seed=7), epsilon=1e-3, normalize=True),      # other layout on stale scratch
        _Device(*_problem_2d(65, seed=11), epsilon=1e-6, normalize=False),            # planar kernel, same array
        _Device(*_problem_3d(1500, 2, 6, seed=2), epsilon=1e-3, normalize=True),      # the first one again
    ]
    expected = []
    for d in problems:
        out = d.outputs()
        need = _run_existing(lib, d, out)
        assert need <= _lib.STATIC_SCRATCH_BYTES
        expected.append(out)
    outs = [d.outputs() for d in problems]            # everything allocated before the calls under test
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    for d, out in zip(problems, outs):                # calls in a row, one shared scratch array
        _run_static(lib, d, out)
    assert torch.cuda.mem_get_info()[0] == free_before, "b200icp_run_static changed the free device memory"
    for want, got in zip(expected, outs):
        _assert_same(want, got)
    # and twice in a row on the same problem
    again = problems[1].outputs()
    _run_static(lib, problems[1], again)
    _assert_same(expected[1], again)


def test_static_scratch_entry_rejects_problems_beyond_its_maximum():
    from lidar_slam_b200 import _lib
    lib = _lib.load()
    d = _Device(*_problem_3d(1500, 2, 2, seed=3), epsilon=1e-3, normalize=True)
    need = ctypes.c_size_t(0)
    big_pairs, big_n = 512, 20000
    assert lib.b200icp_workspace_bytes(big_pairs, big_n, 3, ctypes.byref(need)) == 0
    assert need.value > _lib.STATIC_SCRATCH_BYTES
    out = d.outputs()
    head = d.head()
    head[3], head[9] = big_n, big_pairs               # sizes only: rejected before any pointer is used
    rc = lib.b200icp_run_static(*head, _ptr(out["T"]), _ptr(out["err"]), _ptr(out["iters"]), _ptr(out["status"]),
                                _ptr(out["corr"]), None)
    assert rc == _lib.ERR_STATIC_SCRATCH
    assert b"B200ICP_STATIC_SCRATCH_BYTES" in lib.b200icp_error_string(rc)
    torch.cuda.synchronize()
    assert int((out["status"] == -7).sum()) == out["status"].numel()   # nothing was launched
    with pytest.raises(RuntimeError, match="STATIC_SCRATCH"):
        _lib.check(rc, "b200icp_run_static")
