set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv
python tests/devtools/time_3d.py > gpurun_out/r2a_time3d.log 2>&1
python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/r2a_plain_bench.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:icp2d -s 4 -c 1 -o gpurun_out/r2a_icp2d \
    python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/r2a_ncu_icp2d.log 2>&1
for c in icp3d icp3d_20k; do
  python tests/devtools/profile_cases.py $c > gpurun_out/r2a_plain_$c.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:icp3d -s 1 -c 1 -o gpurun_out/r2a_$c \
      python tests/devtools/profile_cases.py $c > gpurun_out/r2a_ncu_$c.log 2>&1
done
cat gpurun_out/r2a_time3d.log
ls -la gpurun_out
