#!/bin/bash
# compute-sanitizer over a small problem (EAM Cu_u6 6^3, all three force paths): memcheck for
# out-of-bounds / misaligned accesses, racecheck for shared-memory hazards (the block-wide quad counter, the row table and
# the fixed-point accumulators of force_n3.cuh; the chunk table of force_nl.cuh), synccheck for barrier misuse.
# The reference's counterpart is its DEBUG build (Makefile:18-20, CUDA_GET_LAST_ERROR gpu_utility.h:102-133).
# usage (on a GPU box): bash tools/sanitize.sh [out_dir]      -> <out_dir>/r2_sanitizer_{memcheck,racecheck,synccheck}.txt
set -u
OUT=${1:-gpurun_out}
mkdir -p "$OUT"
rc=0
for tool in memcheck racecheck synccheck; do
  # link-cell kernels (TMA-staged tiles), once-per-pair list kernels (shared-memory accumulators), full-list kernels (chunk table);
  # 5 force calls + 12 time steps of EAM Cu_u6 6^3 each, the last ones replayed from the CUDA graph
  for path in "0 " "1 nl_n3=1,graph_step=1" "1 nl_n3=0"; do
    set -- $path
    echo "== compute-sanitizer --tool $tool, COMD_NL=$1 COMD_OPTS=${2:-}" | tee -a "$OUT/r2_sanitizer_$tool.txt"
    COMD_NL=$1 COMD_OPTS=${2:-} timeout 1200 compute-sanitizer --tool $tool --error-exitcode 9 --print-limit 20 \
      python tools/prof_force.py funcfl 6 5 12 2>&1 | grep -v "^$" | tail -12 | tee -a "$OUT/r2_sanitizer_$tool.txt"
    [ ${PIPESTATUS[0]} -ne 0 ] && rc=1
  done
done
exit $rc
