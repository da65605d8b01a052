python -m pytest tests/test_gpu_parity.py tests/test_callers.py -m gpu -x -q 2>&1 | tail -3
echo auto; python tests/devtools/time_3d.py 2>&1 | tail -3 | cut -c1-120
for g in 2 4 8 16; do echo g=$g; B200ICP_CLUSTER_SIZE=$g python tests/devtools/time_3d.py 2>&1 | tail -3 | cut -c1-120; done
