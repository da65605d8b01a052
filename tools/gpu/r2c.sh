set -x
mkdir -p gpurun_out
python tests/devtools/time_variants.py > gpurun_out/r2c_variants.log 2>&1
cat gpurun_out/r2c_variants.log
python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/r2c_plain_bench.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:icp2d_rows -s 4 -c 1 -o gpurun_out/r2c_icp2d_rows \
    python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/r2c_ncu.log 2>&1
ls -la gpurun_out
