#!/bin/bash
# compute-sanitizer passes over the GPU numerics tests (the reference has no
# sanitizer integration for its kernels).  Writes one log per tool under gpurun_out/.
#   benchmarks/sanitize.sh [memcheck|racecheck|synccheck|initcheck] [pytest -k expression]
set -u
TOOL=${1:-memcheck}
EXPR=${2:-"seqpool or table or sparse_pull_push_single or gemm_bf16_tn or ctr_head or cast_transpose or batch_build"}
mkdir -p gpurun_out
LOG=gpurun_out/sanitizer_${TOOL}.log
timeout 1200 compute-sanitizer --tool "$TOOL" --print-limit 20 --error-exitcode 0 \
  python -m pytest tests/test_gpu_kernels.py -q -m gpu -x -p no:cacheprovider --timeout=1000 -k "$EXPR" > "$LOG" 2>&1
echo "exit=$?" >> "$LOG"
grep -E "ERROR SUMMARY|passed|failed|RACECHECK SUMMARY|error" "$LOG" | tail -8
