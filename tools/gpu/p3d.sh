set -x
python -m pytest tests/test_gpu_parity.py tests/test_callers.py tests/test_gpu_fullsize.py -m gpu -x -q 2>&1 | tail -5 > gpurun_out/p3d_pytest.log
python tests/devtools/profile_cases.py icp3d > gpurun_out/p3d_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:icp3d -s 1 -c 1 -o gpurun_out/p3d_prof python tests/devtools/profile_cases.py icp3d > gpurun_out/p3d_ncu.log 2>&1
cat gpurun_out/p3d_pytest.log
