# full GPU suite (with the guard-band tests), bench with the two-stream figure, launch list of the bench command
set -x
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/r2r_pytest.log
cat gpurun_out/r2r_pytest.log
timeout 900 python bench.py --steps 50 > gpurun_out/r2r_bench_n1.json 2> gpurun_out/r2r_bench_n1.err
tail -5 gpurun_out/r2r_bench_n1.err; cut -c1-300 gpurun_out/r2r_bench_n1.json
BENCH="python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-configs"
$BENCH > /dev/null 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 30000 --csv --log-file gpurun_out/r2r_launches.csv \
    $BENCH > gpurun_out/r2r_ncu_launches.log 2>&1
grep -c icp gpurun_out/r2r_launches.csv
