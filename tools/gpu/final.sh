timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 > gpurun_out/r1e_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r1e_smoke.log 2>&1
python bench.py > gpurun_out/r1e_bench_n1.json 2> gpurun_out/r1e_bench_n1.err
python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/r1e_plain_bench.log 2>&1 &&
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/r1e_launches_bench_steps4.csv \
    python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/r1e_ncu_bench.log 2>&1
cat gpurun_out/r1e_pytest.log; tail -1 gpurun_out/r1e_smoke.log; cut -c1-200 gpurun_out/r1e_bench_n1.json; tail -2 gpurun_out/r1e_ncu_bench.log | cut -c1-200
