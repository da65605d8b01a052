# host entries: results written by the kernel straight into page-locked result arrays
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_hostapi.py tests/test_gpu_ranges.py tests/test_gpu_reference_caller.py tests/test_gpu_fullsize.py -x -q 2>&1 | tail -6 > gpurun_out/r2z_pytest.log; cat gpurun_out/r2z_pytest.log
timeout 600 python tests/devtools/time_host_entry.py > gpurun_out/r2z_host_entry.log 2>&1; cat gpurun_out/r2z_host_entry.log
timeout 600 python bench.py --steps 100 --no-configs > gpurun_out/r2z_bench_n1.json 2> gpurun_out/r2z_bench.err; tail -3 gpurun_out/r2z_bench.err; cut -c1-200 gpurun_out/r2z_bench_n1.json
