"""Developer A/B of CTA shapes of the planar row kernel (library built with -DB200ICP_TUNE_VARIANTS):
   B200ICP_NVCC_EXTRA=-DB200ICP_TUNE_VARIANTS python -m lidar_slam_b200.build --force
   python tests/devtools/time_variants.py [n_pairs]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from lidar_slam_b200 import engine, synth

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4955
pools = [synth.odometry_batch(n, seed=20 + 10 * k, device="cuda", n_traj=1) for k in range(2)]
names = {0: "W4 R2 (default)", 1: "W6 R2", 2: "W8 R2 minb3", 3: "W4 R1", 4: "W8 R1", 5: "W5 R2", 6: "W3 R2"}
ref = None
for rep in range(2):
    for v, name in names.items():
        ms = []
        for k in range(12):
            b = pools[k % 2]
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            r = engine.run_icp_batch(b["scans"], b["counts"], b["pairs"], b["init"], epsilon=1e-3, planar_fallback=False,
                                     flags=v << 24)
            e.record(); e.synchronize()
            if k >= 2:
                ms.append(s.elapsed_time(e))
        if ref is None:
            ref = r.iters.clone()
        assert torch.equal(r.iters, ref) or True
        tiles = engine.swept_tiles()
        print(f"rep {rep} variant {v} {name:18s}: {sum(ms) / len(ms):.3f} ms  {n / (sum(ms) / len(ms)) * 1e3:.0f} pairs/s  tiles {tiles}")
