# Developer build of the library with the CTA-shape switch (B200ICP_VARIANT) compiled in; the shipped
# library has no such switch.  Output is git-ignored and travels to the GPU box with the snapshot.
set -e
cd "$(dirname "$0")/../.."
mkdir -p tests/devtools/_tune
cd lidar_slam_b200/csrc
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -shared -Xcompiler -fPIC \
  -DB200ICP_TUNE_VARIANTS ${B200ICP_NVCC_EXTRA} -o ../../tests/devtools/_tune/libb200icp_tune.so \
  api.cu host_api.cu voxel.cu scan_build.cu ogm.cu texture.cu
