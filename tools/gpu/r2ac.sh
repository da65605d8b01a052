# ncu --set full of the row kernel (bench launch) after the instruction-count work
set -x
mkdir -p gpurun_out
BENCH="python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-configs"
$BENCH > /dev/null 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:icp2d_rows -s 4 -c 1 -o gpurun_out/r2ac_icp2d_rows \
    $BENCH > gpurun_out/r2ac_ncu_icp2d.log 2>&1
ls -la gpurun_out/r2ac*
