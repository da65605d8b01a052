python tools/time_host_entry.py 4955 c6 g4 g8 g16 g24 g32 g48 g64 > gpurun_out/gated.log 2>&1
cat gpurun_out/gated.log | grep -v "^\["
timeout 900 python -m pytest tests/test_gpu_hostapi.py tests/test_gpu_prune.py -x -q 2>&1 | tail -3
