# what the driver runs at round end: GPU tests, smoke(), the reference arm, the default bench line
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -6 > gpurun_out/r2f_pytest.log; cat gpurun_out/r2f_pytest.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r2f_smoke.log 2>&1; tail -3 gpurun_out/r2f_smoke.log
timeout 900 python bench.py --impl reference > gpurun_out/r2f_bench_n1_reference_arm.json 2> gpurun_out/r2f_ref.err; cut -c1-300 gpurun_out/r2f_bench_n1_reference_arm.json
SECONDS=0
timeout 1200 python bench.py > gpurun_out/r2f_bench_n1.json 2> gpurun_out/r2f_bench_n1.err
echo "default bench.py wall seconds: $SECONDS" | tee gpurun_out/r2f_bench_wall.txt
tail -3 gpurun_out/r2f_bench_n1.err; cut -c1-300 gpurun_out/r2f_bench_n1.json
