set -x
mkdir -p gpurun_out
timeout 240 python tests/devtools/probe_flow3d.py 8000 12000 16000 20000 > gpurun_out/r2e_probe3d.log 2>&1
cat gpurun_out/r2e_probe3d.log
