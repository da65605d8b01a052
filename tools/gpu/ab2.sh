python tools/tail_probe.py 4955 30000 2>&1 | grep "^n=" | cut -c1-60 > gpurun_out/tail2.log; cat gpurun_out/tail2.log
python tools/prune_ab.py run gpurun_out/ab_pruned.npz > gpurun_out/ab_pruned.log 2>&1
B200ICP_SWEEP=exhaustive python tools/prune_ab.py run gpurun_out/ab_exh.npz > gpurun_out/ab_exh.log 2>&1
python tools/prune_ab.py cmp gpurun_out/ab_pruned.npz gpurun_out/ab_exh.npz | grep -c identical
rm -f gpurun_out/ab_*.npz
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
