python tools/prune_ab.py run /tmp/x.npz > gpurun_out/p2dq_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:icp2d -s 1 -c 1 -o gpurun_out/p2dq_prof python tools/prune_ab.py run /tmp/y.npz > gpurun_out/p2dq_ncu.log 2>&1
tail -2 gpurun_out/p2dq_plain.log | cut -c1-300
