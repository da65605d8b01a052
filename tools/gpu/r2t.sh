# host entry: growing slices, pinned result buffers, cached launch attributes -- tests, slice sweep, bench
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > gpurun_out/r2t_pytest.log
cat gpurun_out/r2t_pytest.log
timeout 600 python tests/devtools/time_host_entry.py > gpurun_out/r2t_host_entry.log 2>&1
cat gpurun_out/r2t_host_entry.log
timeout 900 python bench.py --steps 50 > gpurun_out/r2t_bench_n1.json 2> gpurun_out/r2t_bench_n1.err
tail -5 gpurun_out/r2t_bench_n1.err; cut -c1-200 gpurun_out/r2t_bench_n1.json
