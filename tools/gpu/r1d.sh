set -x
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 > gpurun_out/r1d_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r1d_smoke.log 2>&1
python bench.py > gpurun_out/r1d_bench_n1.json 2> gpurun_out/r1d_bench_n1.err
python bench.py --impl reference --steps 5 > gpurun_out/r1d_bench_n1_reference_arm.json 2> gpurun_out/r1d_ref.err
# launch list of the bench command, after a plain run of the same command
python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/r1d_plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/r1d_launches_bench_steps4.csv \
    python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/r1d_ncu_bench.log 2>&1
# full capture of the dominant kernel (second launch of the same command: a pruned launch of the bench batch)
ncu --set full --clock-control none --import-source on -k regex:icp2d -s 4 -c 1 -o gpurun_out/r1d_icp2d \
    python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/r1d_ncu_full.log 2>&1
cat gpurun_out/r1d_pytest.log; tail -1 gpurun_out/r1d_smoke.log; cut -c1-200 gpurun_out/r1d_bench_n1.json
