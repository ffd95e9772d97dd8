#!/bin/bash
# compute-sanitizer over a small problem (EAM Cu_u6 6^3, all three force paths): memcheck for
# out-of-bounds / misaligned accesses, racecheck for shared-memory hazards (the block-wide quad counter, the row table and
# the fixed-point accumulators of force_n3.cuh; the chunk table of force_nl.cuh), synccheck for barrier misuse.
# The reference's counterpart is its DEBUG build (Makefile:18-20, CUDA_GET_LAST_ERROR gpu_utility.h:102-133).
# usage (on a GPU box): bash tools/sanitize.sh [out_dir]      -> <out_dir>/r2_sanitizer_{memcheck,racecheck,synccheck}.txt
set -u
# The shared B200 pool does not allow compute-sanitizer (profiles/r2_sanitizer_note.txt: runs under it have left GPUs needing
# a reset).  This script therefore refuses to start unless the caller states that the GPU is their own.
if [ "${COMD_SANITIZE_OFF_POOL:-0}" != "1" ]; then
  echo "tools/sanitize.sh: not run. compute-sanitizer is closed on the GPU pool (profiles/r2_sanitizer_note.txt)." >&2
  echo "On a GPU you own, set COMD_SANITIZE_OFF_POOL=1.  On the pool the stand-ins are the bit-identity tests across" >&2
  echo "scheduling choices (tests/test_gpu_parity.py) and the ERRB_* guards of the force kernels." >&2
  exit 3
fi
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
