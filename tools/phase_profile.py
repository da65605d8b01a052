"""Developer tool: cycles per phase of the planar ICP kernel (needs a library built with
B200ICP_NVCC_EXTRA=-DB200ICP_PHASE_PROFILE; counters live in bytes 64..127 of the workspace)."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from lidar_slam_b200 import _lib, synth
from lidar_slam_b200.engine import _ptr, _stream_ptr

lib = _lib.load()
n = 4955
b = synth.odometry_batch(n, seed=20, device="cuda")
scans, counts, pairs, init = b["scans"], b["counts"], b["pairs"], b["init"]
eps = torch.full((n,), 1e-3, dtype=torch.float64, device="cuda")
norm = torch.zeros((n,), dtype=torch.uint8, device="cuda")
T = torch.empty((n, 4, 4), dtype=torch.float64, device="cuda")
err = torch.empty((n,), dtype=torch.float64, device="cuda")
it = torch.empty((n,), dtype=torch.int32, device="cuda")
st = torch.empty((n,), dtype=torch.int32, device="cuda")
nb = ctypes.c_size_t()
lib.b200icp_workspace_bytes(n, scans.shape[1], 2, ctypes.byref(nb))
ws = torch.zeros((nb.value,), dtype=torch.uint8, device="cuda")
for rep in range(3):
    _lib.check(lib.b200icp_run(_ptr(scans), _ptr(counts), scans.shape[0], scans.shape[1], 2, _ptr(pairs), _ptr(init),
                               _ptr(eps), _ptr(norm), n, 2000, 1e-4, _ptr(T), _ptr(err), _ptr(it), _ptr(st), None,
                               _ptr(ws), ws.numel(), _stream_ptr()), "run")
torch.cuda.synchronize()
acc = ws[64:128].view(torch.int64).cpu().numpy().astype(np.float64)
names = ["stage+stats", "transform+sweep", "shared-row barrier+merge", "resolve", "sums+warp reduce",
         "exchange barrier+totals", "solve+stop", "pop/finalise"]
tot = acc.sum()
passes = float(it.sum().item())
print(f"total passes {passes:.0f}; cycles of warp 0 thread 0 summed over CTAs: {tot:.3e}")
for k, nm in enumerate(names):
    print(f"  {nm:28s} {100 * acc[k] / tot:5.1f} %   {acc[k] / passes:9.0f} cycles per pass")
