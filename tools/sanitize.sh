#!/bin/bash
# compute-sanitizer over the small smoke problem (EAM Cu_u6 6^3, both force paths, 5 steps each): memcheck for
# out-of-bounds / misaligned accesses, racecheck for shared-memory hazards (the block-wide quad counter, the row table and
# the fixed-point accumulators of force_n3.cuh; the chunk table of force_nl.cuh), synccheck for barrier misuse.
# The reference's counterpart is its DEBUG build (Makefile:18-20, CUDA_GET_LAST_ERROR gpu_utility.h:102-133).
# usage (on a GPU box): bash tools/sanitize.sh [out_dir]      -> <out_dir>/r2_sanitizer_{memcheck,racecheck,synccheck}.txt
set -u
OUT=${1:-gpurun_out}
mkdir -p "$OUT"
rc=0
for tool in memcheck racecheck synccheck; do
  for n3 in 1 0; do
    echo "== compute-sanitizer --tool $tool, COMD_NL_N3=$n3" | tee -a "$OUT/r2_sanitizer_$tool.txt"
    COMD_NL_N3=$n3 timeout 1200 compute-sanitizer --tool $tool --error-exitcode 9 --print-limit 20 \
      python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | grep -v "^$" | tail -25 | tee -a "$OUT/r2_sanitizer_$tool.txt"
    [ ${PIPESTATUS[0]} -ne 0 ] && rc=1
  done
done
exit $rc
