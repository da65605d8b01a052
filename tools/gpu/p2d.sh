python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/p2d_plain.log 2>&1 && timeout 300 ncu --set full --clock-control none --import-source on -k regex:icp2d -s 4 -c 1 -o gpurun_out/r1d_icp2d_final python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/p2d_ncu.log 2>&1
tail -2 gpurun_out/p2d_ncu.log | cut -c1-200
