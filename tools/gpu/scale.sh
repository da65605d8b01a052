N=$1
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/scale_n$N.json 2> gpurun_out/scale_n$N.err
cut -c1-220 gpurun_out/scale_n$N.json
if [ "$N" = "8" ]; then
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 5 --warmup 3 --pairs 125000 --no-cpu-baseline > gpurun_out/scale_n${N}_config5.json 2> gpurun_out/scale_n${N}_config5.err
cut -c1-220 gpurun_out/scale_n${N}_config5.json
python -m pytest tests/test_gpu_multi.py -m gpu -x -q 2>&1 | tail -3
fi
