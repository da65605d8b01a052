for v in 0 1 2 3 4 5; do
  echo "variant $v"; B200ICP_VARIANT=$v python tools/tail_probe.py 4955 30000 2>&1 | grep "^n=" | cut -c1-60
done > gpurun_out/variants.log 2>&1
cat gpurun_out/variants.log
python tools/prune_ab.py run gpurun_out/ab_pruned.npz > gpurun_out/ab_pruned.log 2>&1
B200ICP_SWEEP=exhaustive python tools/prune_ab.py run gpurun_out/ab_exh.npz > gpurun_out/ab_exh.log 2>&1
python tools/prune_ab.py cmp gpurun_out/ab_pruned.npz gpurun_out/ab_exh.npz | grep -c identical
rm -f gpurun_out/ab_*.npz
