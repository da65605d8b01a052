# dataflow kernel: one query per lane (rows of 32, items of 256 points) and / or 3 CTAs per SM against the shipped 2 x 2
set -x
mkdir -p gpurun_out
for lib in lidar_slam_b200/libb200icp.so tests/devtools/_tune/libb200icp_flowq1.so tests/devtools/_tune/libb200icp_flowq1m3.so tests/devtools/_tune/libb200icp_flowq2m3.so; do
  echo "== $lib"
  B200ICP_LIB=$lib timeout 300 python tests/devtools/time_3d.py 2>&1 | tail -3 | cut -c1-150
  B200ICP_LIB=$lib timeout 300 python tests/devtools/time_3d.py 2>&1 | tail -3 | cut -c1-150
done > gpurun_out/r2y_flow_q.log 2>&1
cat gpurun_out/r2y_flow_q.log
B200ICP_LIB=tests/devtools/_tune/libb200icp_flowq1.so timeout 300 python -m pytest tests/test_gpu_flow3d.py -x -q 2>&1 | tail -3
