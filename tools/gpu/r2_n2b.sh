# 2 GPUs: per-rank kernel times (different trajectories per rank) and the same run with rank 0's data on both ranks
set -x
mkdir -p gpurun_out
run() { tag=$1; shift
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29502 \
    bench.py --gpus 2 --steps 100 --warmup 3 --no-cpu-baseline --no-configs "$@" > gpurun_out/r2c_bench_n2$tag.json 2> gpurun_out/r2c_bench_n2$tag.err
  tail -2 gpurun_out/r2c_bench_n2$tag.err; cut -c1-200 gpurun_out/r2c_bench_n2$tag.json; }
run ""
run _same_data --same-data
timeout 300 python bench.py --steps 100 --no-cpu-baseline --no-configs > gpurun_out/r2c_bench_n1_same_box.json 2>/dev/null
cut -c1-200 gpurun_out/r2c_bench_n1_same_box.json
