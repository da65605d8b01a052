# batched range loads in the kernel's scan construction; equal slices again; upload rate
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_ranges.py tests/test_gpu_hostapi.py tests/test_gpu_canaries.py tests/test_gpu_fullsize.py -x -q 2>&1 | tail -8 > gpurun_out/r2u_pytest.log
cat gpurun_out/r2u_pytest.log
timeout 600 python tests/devtools/time_host_entry.py > gpurun_out/r2u_host_entry.log 2>&1
cat gpurun_out/r2u_host_entry.log
timeout 900 python bench.py --steps 50 > gpurun_out/r2u_bench_n1.json 2> gpurun_out/r2u_bench_n1.err
tail -5 gpurun_out/r2u_bench_n1.err; cut -c1-200 gpurun_out/r2u_bench_n1.json
