# row kernel: per-lane admission of the blocks that do not overlap the row box
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_prune.py tests/test_gpu_parity.py tests/test_gpu_fullsize.py tests/test_gpu_ranges.py -x -q 2>&1 | tail -6 > gpurun_out/r2aa_pytest.log; cat gpurun_out/r2aa_pytest.log
timeout 600 python bench.py --steps 100 --no-configs --no-cpu-baseline > gpurun_out/r2aa_bench_n1.json 2> gpurun_out/r2aa_bench.err; tail -3 gpurun_out/r2aa_bench.err; cut -c1-200 gpurun_out/r2aa_bench_n1.json
