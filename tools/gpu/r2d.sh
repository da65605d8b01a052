set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_flow3d.py tests/test_gpu_parity.py tests/test_callers.py -m gpu -x -q 2>&1 | tail -25 > gpurun_out/r2d_pytest3d.log
cat gpurun_out/r2d_pytest3d.log
timeout 300 python tests/devtools/time_3d.py > gpurun_out/r2d_time3d.log 2>&1
cat gpurun_out/r2d_time3d.log
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > gpurun_out/r2d_pytest_all.log
cat gpurun_out/r2d_pytest_all.log
