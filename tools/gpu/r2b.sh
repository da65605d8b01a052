set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/r2b_pytest.log
cat gpurun_out/r2b_pytest.log
python bench.py --steps 50 --no-cpu-baseline > gpurun_out/r2b_bench_n1.json 2> gpurun_out/r2b_bench_n1.err
tail -3 gpurun_out/r2b_bench_n1.err; cut -c1-400 gpurun_out/r2b_bench_n1.json
