"""Where the end-to-end step of bench.py (b200icp_run_host_ranges, one gated launch under the upload)
spends its time: the same call with different numbers of pool slices, against the resident kernel.
   python tests/devtools/time_host_entry.py [n_pairs]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
from lidar_slam_b200 import engine, hostapi, synth, _lib

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4955
b = synth.odometry_batch(n, seed=20, device="cpu", n_traj=1, with_ranges=True)
pin = lambda a: torch.as_tensor(np.ascontiguousarray(a)).pin_memory().numpy()
ranges, pairs, init = pin(b["ranges"].numpy()), pin(b["pairs"].numpy().astype(np.int32)), pin(b["init"].numpy())
ws = hostapi.HostRangesWorkspace(n, ranges.shape[0], ranges.shape[1] + (ranges.shape[1] & 1), False)
tables = hostapi.beam_tables(ranges.shape[1])
dev_r = torch.as_tensor(ranges).cuda()
for _ in range(3):
    engine.run_icp_ranges_batch(dev_r, torch.as_tensor(pairs).cuda(), torch.as_tensor(init).cuda(), epsilon=1e-3)
torch.cuda.synchronize()
ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
ev[0].record()
for _ in range(10):
    engine.run_icp_ranges_batch(dev_r, torch.as_tensor(pairs).cuda(), torch.as_tensor(init).cuda(), epsilon=1e-3)
ev[1].record(); torch.cuda.synchronize()
print(f"resident ranges kernel (+ small uploads): {ev[0].elapsed_time(ev[1]) / 10:.3f} ms")
# the upload alone: one pinned -> device copy of the range rows
src_t, dst_t = torch.as_tensor(ranges), torch.empty_like(dev_r)
for _ in range(3):
    dst_t.copy_(src_t, non_blocking=True)
torch.cuda.synchronize()
ev[0].record()
for _ in range(10):
    dst_t.copy_(src_t, non_blocking=True)
ev[1].record(); torch.cuda.synchronize()
ms_copy = ev[0].elapsed_time(ev[1]) / 10
print(f"upload alone: {ranges.nbytes / 1e6:.1f} MB in {ms_copy:.3f} ms = {ranges.nbytes / ms_copy / 1e6:.1f} GB/s")
for slices, direct in ((0, True), (0, False), (8, True), (16, True), (16, False), (24, True), (32, True)):
    fl = (_lib.flag_slices(slices) if slices else 0) | (0 if direct else _lib.FLAG_HOST_NO_DIRECT)
    best, tot = 1e9, 0.0
    for rep in range(14):
        t0 = time.perf_counter()
        hostapi.run_icp_ranges_host(ranges, pairs, init, epsilon=1e-3, flags=fl, out=ws, tables=tables)
        dt = (time.perf_counter() - t0) * 1e3
        if rep >= 4:
            best, tot = min(best, dt), tot + dt
    print(f"slices {slices or 'default'}{'' if direct else ' (results copied back)'}: best {best:.3f} ms  mean {tot / 10:.3f} ms  {n / (tot / 10) * 1e3:.0f} pairs/s")
