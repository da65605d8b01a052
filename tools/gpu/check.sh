python -m pytest tests -m gpu -x -q 2>&1 | tail -4 > gpurun_out/c_pytest.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/c_smoke.log 2>&1
python bench.py > gpurun_out/c_bench.json 2> gpurun_out/c_bench.err
python bench.py --impl reference > gpurun_out/c_ref.json 2> gpurun_out/c_ref.err
cat gpurun_out/c_pytest.log; tail -2 gpurun_out/c_smoke.log; cut -c1-400 gpurun_out/c_bench.json; cut -c1-300 gpurun_out/c_ref.json
