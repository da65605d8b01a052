set -x
B200ICP_NVCC_EXTRA=-DB200ICP_PHASE_PROFILE python -c "from lidar_slam_b200 import build; build.build(force=True)"
python tools/phase_profile.py > gpurun_out/phase_pruned.log 2>&1
tail -12 gpurun_out/phase_pruned.log
