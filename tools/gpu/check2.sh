timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > gpurun_out/c2_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/c2_smoke.log 2>&1
python bench.py > gpurun_out/c2_bench.json 2> gpurun_out/c2_bench.err
python bench.py --impl reference --steps 5 > gpurun_out/c2_ref.json 2> gpurun_out/c2_ref.err
cat gpurun_out/c2_pytest.log; tail -2 gpurun_out/c2_smoke.log; cut -c1-300 gpurun_out/c2_bench.json; cut -c1-200 gpurun_out/c2_ref.json
