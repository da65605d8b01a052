set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/r2i_pytest.log
cat gpurun_out/r2i_pytest.log
timeout 900 python bench.py --steps 50 > gpurun_out/r2i_bench_n1.json 2> gpurun_out/r2i_bench_n1.err
tail -5 gpurun_out/r2i_bench_n1.err; cut -c1-200 gpurun_out/r2i_bench_n1.json
for i in 1 2 3; do timeout 300 python tests/devtools/time_3d.py 2>&1 | tail -3; done > gpurun_out/r2i_time3d.log
cat gpurun_out/r2i_time3d.log
