# Profiles of round 2: launch list of the bench command, ncu --set full of the row kernel (bench
# launch) and of the dataflow kernel (6 000- and 20 000-point yaw sweeps).  Each ncu run only after the
# same command exited 0 without it.
set -x
mkdir -p gpurun_out
BENCH="python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-configs"
$BENCH > gpurun_out/r2q_plain_bench.json 2> gpurun_out/r2q_plain_bench.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2q_launches.csv \
    $BENCH > gpurun_out/r2q_ncu_launches.log 2>&1
$BENCH > /dev/null 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:icp2d_rows -s 4 -c 1 -o gpurun_out/r2q_icp2d_rows \
    $BENCH > gpurun_out/r2q_ncu_icp2d.log 2>&1
for c in icp3d icp3d_20k; do
  python tests/devtools/profile_cases.py $c > gpurun_out/r2q_plain_$c.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:icp3d_flow -s 1 -c 1 -o gpurun_out/r2q_$c \
      python tests/devtools/profile_cases.py $c > gpurun_out/r2q_ncu_$c.log 2>&1
done
ls -la gpurun_out
