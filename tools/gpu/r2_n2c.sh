# 2 GPUs: pairs retired once per CTA (one system fence per CTA), submit/wait host entries; tests + N = 2 lines
set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_multi.py tests/test_gpu_ranges.py tests/test_gpu_hostapi.py -x -q 2>&1 | tail -8 > gpurun_out/r2d_pytest.log
cat gpurun_out/r2d_pytest.log
run() { tag=$1; shift
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29502 \
    bench.py --gpus 2 --steps 100 --warmup 3 --no-cpu-baseline --no-configs "$@" > gpurun_out/r2d_bench_n2$tag.json 2> gpurun_out/r2d_bench_n2$tag.err
  tail -2 gpurun_out/r2d_bench_n2$tag.err; cut -c1-200 gpurun_out/r2d_bench_n2$tag.json; }
run ""
run _same_data --same-data
timeout 300 python bench.py --steps 100 --no-cpu-baseline --no-configs > gpurun_out/r2d_bench_n1_same_box.json 2> gpurun_out/r2d_bench_n1.err
tail -3 gpurun_out/r2d_bench_n1.err; cut -c1-200 gpurun_out/r2d_bench_n1_same_box.json
