TOOL=${1:-memcheck}
timeout 120 python bench/sanitize_once.py || exit 1
timeout 500 compute-sanitizer --tool $TOOL --print-limit 20 python bench/sanitize_once.py > gpurun_out/sanitizer_$TOOL.log 2>&1; echo "compute-sanitizer $TOOL exit $?"
grep -v "^$" gpurun_out/sanitizer_$TOOL.log | tail -12
