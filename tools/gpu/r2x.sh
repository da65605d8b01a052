# row kernel with the source read from the pool (5 or 6 CTAs/SM) against the shipped shape
set -x
mkdir -p gpurun_out
B200ICP_LIB=tests/devtools/_tune/libb200icp_tune.so timeout 600 python tests/devtools/time_variants.py > gpurun_out/r2x_variants.log 2>&1; grep -E "variant (0|9|10|11) " gpurun_out/r2x_variants.log; tail -3 gpurun_out/r2x_variants.log
