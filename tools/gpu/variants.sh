for v in 0 1 2 3; do
  echo "variant $v"; B200ICP_VARIANT=$v python tools/tail_probe.py 4955 30000 2>&1 | grep "^n=" | cut -c1-60
done > gpurun_out/variants2.log 2>&1
cat gpurun_out/variants2.log
