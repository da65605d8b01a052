set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -q 2>&1 | tail -15 > gpurun_out/r2p_noalloc_alone.log
cat gpurun_out/r2p_noalloc_alone.log
timeout 1800 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/r2p_pytest.log
cat gpurun_out/r2p_pytest.log
timeout 900 python bench.py --steps 50 > gpurun_out/r2p_bench_n1.json 2> gpurun_out/r2p_bench_n1.err
tail -5 gpurun_out/r2p_bench_n1.err; cut -c1-300 gpurun_out/r2p_bench_n1.json
B200ICP_LIB=tests/devtools/_tune/libb200icp_tune.so timeout 600 python tests/devtools/time_variants.py > gpurun_out/r2p_variants.log 2>&1; tail -20 gpurun_out/r2p_variants.log
