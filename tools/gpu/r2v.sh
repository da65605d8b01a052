# pass-count predictor probe
set -x
mkdir -p gpurun_out
timeout 600 python tests/devtools/probe_pass_predictor.py > gpurun_out/r2v_pass_predictor.log 2>&1
cat gpurun_out/r2v_pass_predictor.log
