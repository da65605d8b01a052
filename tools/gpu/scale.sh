N=$1
if [ "$N" = "2" ]; then
timeout 300 python -m pytest tests/test_gpu_multi.py -m gpu -x -q 2>&1 | tail -3
fi
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 100 --warmup 3 --no-cpu-baseline > gpurun_out/r1d_bench_n$N.json 2> gpurun_out/r1d_bench_n$N.err
cut -c1-220 gpurun_out/r1d_bench_n$N.json; tail -2 gpurun_out/r1d_bench_n$N.err
if [ "$N" = "8" ]; then
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 5 --warmup 3 --pairs 125000 --no-cpu-baseline > gpurun_out/r1d_bench_n8_config5_1M_pairs.json 2> gpurun_out/r1d_bench_n8_config5.err
cut -c1-220 gpurun_out/r1d_bench_n8_config5_1M_pairs.json
fi
