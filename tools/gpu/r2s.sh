# dataflow kernel with decoupled warps: 3-D tests first (under a timeout: a protocol bug would hang), timings, host-entry slices
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_flow3d.py tests/test_gpu_canaries.py tests/test_zz_no_alloc.py -x -q 2>&1 | tail -15 > gpurun_out/r2s_pytest3d.log
cat gpurun_out/r2s_pytest3d.log
for i in 1 2; do timeout 300 python tests/devtools/time_3d.py 2>&1 | tail -4; done > gpurun_out/r2s_time3d.log
cat gpurun_out/r2s_time3d.log
timeout 600 python tests/devtools/time_host_entry.py > gpurun_out/r2s_host_entry.log 2>&1
cat gpurun_out/r2s_host_entry.log
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -6 > gpurun_out/r2s_pytest_all.log
cat gpurun_out/r2s_pytest_all.log
