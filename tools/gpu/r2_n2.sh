# 2 GPUs: exchange tests + bench at N = 2 (PeerGather with 4 row slots, wait deferred by one step), short timeouts
set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_multi.py -x -q 2>&1 | tail -8 > gpurun_out/r2b_n2_pytest_multi.log
cat gpurun_out/r2b_n2_pytest_multi.log
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29502 \
    bench.py --gpus 2 --steps 100 --warmup 3 --no-cpu-baseline > gpurun_out/r2b_bench_n2.json 2> gpurun_out/r2b_bench_n2.err
tail -5 gpurun_out/r2b_bench_n2.err; cut -c1-250 gpurun_out/r2b_bench_n2.json
timeout 300 python bench.py --steps 100 --no-cpu-baseline --no-configs > gpurun_out/r2b_bench_n1_same_box.json 2>/dev/null
cut -c1-250 gpurun_out/r2b_bench_n1_same_box.json
