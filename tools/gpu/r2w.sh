# randomised parity of the final round-2 kernels against the oracle (row kernel, dataflow kernel)
set -x
mkdir -p gpurun_out
timeout 500 python tests/devtools/fuzz_parity.py 240 21 > gpurun_out/r2w_fuzz_small.log 2>&1; tail -3 gpurun_out/r2w_fuzz_small.log
timeout 400 python tests/devtools/fuzz_parity.py 150 22 big > gpurun_out/r2w_fuzz_big.log 2>&1; tail -3 gpurun_out/r2w_fuzz_big.log
