# weak scaling on one 8-GPU box: the 2-GPU tests, then bench at N = 8, 4, 2 launched the way the driver does
set -x
mkdir -p gpurun_out
nvidia-smi -L | wc -l
timeout 300 python -m pytest tests/test_gpu_multi.py -x -q 2>&1 | tail -5 > gpurun_out/r2g_scale_pytest_multi.log
cat gpurun_out/r2g_scale_pytest_multi.log
run() { N=$1; shift
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + N)) \
    bench.py --gpus $N --steps 100 --warmup 3 "$@" > gpurun_out/r2g_bench_n$N.json 2> gpurun_out/r2g_bench_n$N.err
  tail -3 gpurun_out/r2g_bench_n$N.err; cut -c1-250 gpurun_out/r2g_bench_n$N.json; }
run 8 --no-cpu-baseline
run 4 --no-cpu-baseline --no-configs
run 2 --no-cpu-baseline --no-configs
timeout 300 python bench.py --steps 100 --no-cpu-baseline --no-configs > gpurun_out/r2g_bench_n1_same_box.json 2>/dev/null
cut -c1-250 gpurun_out/r2g_bench_n1_same_box.json
