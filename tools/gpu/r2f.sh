set -x
mkdir -p gpurun_out
for n in 9000 12000; do
for m in noprune prune nn; do
timeout 40 python tests/devtools/probe_flow3d_b.py $n $m >> gpurun_out/r2f_probe.log 2>&1; echo "exit $? ($n $m)" >> gpurun_out/r2f_probe.log
done; done
cat gpurun_out/r2f_probe.log
