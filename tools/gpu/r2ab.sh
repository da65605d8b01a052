# row kernel: rows cut on a grid aligned with the target blocks (index shift from the previous pass)
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > gpurun_out/r2ab_pytest.log; cat gpurun_out/r2ab_pytest.log
timeout 600 python bench.py --steps 100 --no-cpu-baseline > gpurun_out/r2ab_bench_n1.json 2> gpurun_out/r2ab_bench.err; tail -3 gpurun_out/r2ab_bench.err; cut -c1-200 gpurun_out/r2ab_bench_n1.json
