python bench.py --steps 20 --warmup 3 > gpurun_out/b1_bench.json 2> gpurun_out/b1_bench.err
cut -c1-1500 gpurun_out/b1_bench.json; tail -3 gpurun_out/b1_bench.err
python tools/time_host_entry.py 4955 1 2 3 4 6 8 12 16 > gpurun_out/b1_chunks.log 2>&1
cat gpurun_out/b1_chunks.log
