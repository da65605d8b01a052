python tools/time_host_entry.py 4955 c6 g8 g12 g16 g24 g32 > gpurun_out/gated.log 2>&1
cat gpurun_out/gated.log | grep -v "^\["
timeout 900 python -m pytest tests/test_gpu_hostapi.py tests/test_gpu_prune.py tests/test_gpu_parity.py -x -q 2>&1 | tail -3
python bench.py --steps 50 --no-cpu-baseline > gpurun_out/g_bench.json 2>gpurun_out/g_bench.err; python -c "
import json; d=json.load(open('gpurun_out/g_bench.json')); print(d['value'], d['e2e']['value'], d['e2e']['torch_engine_path']['value'])"
# sanity: the gated paths under a tool that serialises launches must not stall
time (CUDA_LAUNCH_BLOCKING=1 python tools/time_host_entry.py 4955 g24 2>&1 | tail -1)
