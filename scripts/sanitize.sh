#!/bin/bash
# compute-sanitizer over the kernel tests (race / memory checking for the new concurrency).
# ONE tool per invocation (B200_PROFILING.md): scripts/sanitize.sh memcheck|racecheck|synccheck|initcheck [pytest -k expr]
TOOL=${1:-memcheck}
EXPR=${2:-"adamw or mix_n or colsum or layernorm or embed or head or sha256 or block_validate or pagerank"}
OUT=gpurun_out/sanitizer_${TOOL}.log
mkdir -p gpurun_out
timeout 900 compute-sanitizer --tool "$TOOL" --error-exitcode 99 --print-limit 20 \
    python -m pytest tests/test_gpu_kernels.py tests/test_gpu_fed.py -m gpu -q -x -k "$EXPR" -p no:cacheprovider > "$OUT" 2>&1
RC=$?
grep -E "ERROR SUMMARY|passed|failed|error" "$OUT" | tail -5
echo "sanitizer[$TOOL] exit code $RC"
exit $RC
