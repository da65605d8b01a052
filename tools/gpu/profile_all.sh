set -x
mkdir -p gpurun_out
# launch list of the bench command (shares of the step), after a plain run of the same command
python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/p_plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/p_launches.csv \
    python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/p_ncu_bench.log 2>&1
for c in ogm texture icp3d scans; do
  case $c in ogm) k=ogm_apply;; texture) k=texture_paint;; icp3d) k=icp3d;; scans) k=ranges_to_scans;; esac
  python tests/devtools/profile_cases.py $c > gpurun_out/p_plain_$c.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:$k -s 1 -c 1 -o gpurun_out/p_$c \
      python tests/devtools/profile_cases.py $c > gpurun_out/p_ncu_$c.log 2>&1
done
ls -la gpurun_out | head -40
