"""Developer A/B of CTA shapes of the planar row kernel.  Needs a library built with
-DB200ICP_TUNE_VARIANTS (tools/gpu/build_tune_lib.sh writes tests/devtools/_tune/libb200icp_tune.so):
   B200ICP_LIB=tests/devtools/_tune/libb200icp_tune.so python tests/devtools/time_variants.py [n_pairs]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from lidar_slam_b200 import engine, synth

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4955
pools = [synth.odometry_batch(n, seed=20 + 10 * k, device="cuda", n_traj=1) for k in range(2)]
names = {0: "W4 R1 minb4 (default)", 1: "W4 R2 minb4", 2: "W8 R2 minb2", 3: "W6 R1 minb3", 4: "W8 R1 minb2",
         5: "W5 R1 minb3", 6: "W3 R1 minb5", 7: "W4 R3 minb4", 8: "W2 R1 minb8",
         9: "W4 R1 minb5 source in global", 10: "W4 R1 minb4 source in global", 11: "W4 R1 minb6 source in global"}
ref = None
for rep in range(2):
    for v, name in names.items():
        ms = []
        for k in range(12):
            b = pools[k % 2]
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            os.environ["B200ICP_VARIANT"] = str(v)
            r = engine.run_icp_batch(b["scans"], b["counts"], b["pairs"], b["init"], epsilon=1e-3, planar_fallback=False)
            e.record(); e.synchronize()
            if k >= 2:
                ms.append(s.elapsed_time(e))
        if ref is None:
            ref = r.iters.clone()
        if v in (9, 10, 11):
            assert torch.equal(r.iters, ref) and torch.equal(r.T, ref_T), name   # same arithmetic and order: same bits
        if v == 0:
            ref_T = r.T.clone()
        tiles = engine.swept_tiles()
        print(f"rep {rep} variant {v} {name:18s}: {sum(ms) / len(ms):.3f} ms  {n / (sum(ms) / len(ms)) * 1e3:.0f} pairs/s  tiles {tiles}")
