# compute-sanitizer over tests/devtools/sanitize_cases.py, ONE tool per gpurun call:
#   gpurun --timeout 2400 -- 'bash tools/gpu/r2_sanitize.sh memcheck'   (racecheck, synccheck)
# The full log is copied to profiles/r2_sanitizer_<tool>.txt by hand afterwards.
set -x
TOOL=${1:-memcheck}
mkdir -p gpurun_out
python tests/devtools/sanitize_cases.py > gpurun_out/r2_sanitize_plain.log 2>&1 || { tail -20 gpurun_out/r2_sanitize_plain.log; exit 1; }
timeout 2000 compute-sanitizer --tool $TOOL --print-limit 50 python tests/devtools/sanitize_cases.py > gpurun_out/r2_sanitizer_$TOOL.txt 2>&1
echo "exit code $?" >> gpurun_out/r2_sanitizer_$TOOL.txt
tail -15 gpurun_out/r2_sanitizer_$TOOL.txt
