# the default bench line (200 steps), wall clock around it
set -x
mkdir -p gpurun_out
start=$(date +%s.%N)
timeout 1200 python bench.py > gpurun_out/r2f_bench_n1.json 2> gpurun_out/r2f_bench_n1.err
end=$(date +%s.%N); echo "bench wall seconds: $(echo "$end - $start" | bc)" | tee gpurun_out/r2f_bench_wall.txt
tail -3 gpurun_out/r2f_bench_n1.err; cut -c1-300 gpurun_out/r2f_bench_n1.json
