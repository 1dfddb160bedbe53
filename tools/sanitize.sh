#!/bin/bash
# compute-sanitizer over the GPU tests that exercise the multi-stream machinery (side / ingest / pose streams, tickets,
# the persistent RANSAC kernel, speculative steps and their rollbacks):  gpurun --timeout 2400 -- 'bash tools/sanitize.sh a'
TAG=${1:-a}
O=gpurun_out
mkdir -p $O
SEL='pipelined or rollback or early or prefetch or frame_ahead or append or speculat or pnp_vs_golden or klt_vs_golden or promotion_step'
for TOOL in memcheck racecheck initcheck synccheck; do
  timeout 900 compute-sanitizer --tool $TOOL --error-exitcode 9 --print-limit 20 \
      python -m pytest tests -m gpu -x -q -k "$SEL" > $O/r2_sanitizer_${TOOL}_$TAG.log 2>&1
  echo "$TOOL rc=$?"; grep -E "ERROR SUMMARY|passed|failed|error" $O/r2_sanitizer_${TOOL}_$TAG.log | tail -4
done
