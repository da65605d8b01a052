set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/r2h_pytest.log
cat gpurun_out/r2h_pytest.log
timeout 300 python tests/devtools/time_3d.py > gpurun_out/r2h_time3d.log 2>&1
cat gpurun_out/r2h_time3d.log
timeout 900 python bench.py > gpurun_out/r2h_bench_n1.json 2> gpurun_out/r2h_bench_n1.err
tail -5 gpurun_out/r2h_bench_n1.err; cut -c1-300 gpurun_out/r2h_bench_n1.json
timeout 600 python bench.py --impl reference --steps 5 > gpurun_out/r2h_bench_ref.json 2> gpurun_out/r2h_bench_ref.err
cut -c1-300 gpurun_out/r2h_bench_ref.json
