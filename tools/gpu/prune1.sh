set -x
python tools/prune_ab.py run gpurun_out/ab_pruned.npz > gpurun_out/ab_pruned.log 2>&1
B200ICP_SWEEP=exhaustive python tools/prune_ab.py run gpurun_out/ab_exh.npz > gpurun_out/ab_exh.log 2>&1
python tools/prune_ab.py cmp gpurun_out/ab_pruned.npz gpurun_out/ab_exh.npz > gpurun_out/ab_cmp.log 2>&1
tail -3 gpurun_out/ab_pruned.log gpurun_out/ab_exh.log; cat gpurun_out/ab_cmp.log
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/p1_pytest.log
cat gpurun_out/p1_pytest.log
rm -f gpurun_out/ab_*.npz
