"""Developer timing of the host-buffer entry b200icp_run_host on the config-2 batch
(B200ICP_PIPELINE_CHUNKS is read at every call)."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from lidar_slam_b200 import hostapi, synth

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4955
b = synth.odometry_batch(n, seed=20, device="cuda", n_traj=1)
hp = {k: b[k].cpu().pin_memory().numpy() for k in ("scans", "counts", "pairs", "init")}
ref = None
for mode in sys.argv[2:] or ["g24", "c6"]:          # gN: gated launch, N pool slices; cN: N chunked launches
    if mode[0] == "g":
        os.environ["B200ICP_HOST_GATED"] = "1"
        os.environ["B200ICP_HOST_SLICES"] = mode[1:]
    else:
        os.environ["B200ICP_HOST_GATED"] = "0"
        os.environ["B200ICP_PIPELINE_CHUNKS"] = mode[1:]
    best = 1e9
    for rep in range(10):
        t0 = time.perf_counter()
        res = hostapi.run_icp_batch_host(hp["scans"], hp["counts"], hp["pairs"], hp["init"], epsilon=1e-3)
        dt = time.perf_counter() - t0
        if rep >= 2:
            best = min(best, dt)
    if ref is None:
        ref = res
    same = all((getattr(ref, k) == getattr(res, k)).all() for k in ("T", "error", "iters", "status"))
    print(f"{mode}: best {best * 1e3:.3f} ms  {n / best:.0f} pairs/s  ({hp['scans'].nbytes / best / 1e9:.1f} GB/s of pool)"
          f"  same results as {sys.argv[2] if len(sys.argv) > 2 else 'g24'}: {same}")
