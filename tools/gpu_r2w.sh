#!/bin/bash
# round 2: link-cell EAM path with the per-step hit list (tile_hits): tests, then A/B on C3 and C5
set -u
mkdir -p gpurun_out
T=r2w
python -m pytest tests -m gpu -q -x > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?"
grep -E "^E  |^(FAILED|ERROR)|passed|failed" gpurun_out/${T}_pytest.log | cut -c1-300 | head -30
export COMD_NL=0
for sz in "setfl 60" "funcfl 40" "setfl 160"; do for o in "tile_hits=0" "tile_hits=1"; do
  COMD_OPTS=$o timeout 600 python tools/prof_force.py $sz 3 20 2>&1 | grep -E "step|force " | sed "s/^/[$sz $o] /"
done; done | tee gpurun_out/${T}_tile_hits_ab.txt
