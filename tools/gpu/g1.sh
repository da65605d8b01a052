set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/s1_pytest.log
python bench.py --steps 10 --warmup 3 > gpurun_out/s1_bench.json 2> gpurun_out/s1_bench.err
python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/s1_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:icp2d -s 4 -c 1 -o gpurun_out/s1_prof python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/s1_ncu.log 2>&1
ls -la gpurun_out
