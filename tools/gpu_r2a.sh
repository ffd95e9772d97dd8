#!/bin/bash
# round 2, first contact of the once-per-pair list kernels: GPU tests, then A/B against the full-list kernels
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -25 | tee gpurun_out/r2a_pytest.txt
export COMD_NL=1
for o in "nl_n3=0" "" "nl_block=40402" "nl_block=40404,nl_threads=384" "nl_block=80404"; do
  for pot in setfl lj; do
    COMD_OPTS=$o timeout 300 python tools/prof_force.py $pot 60 5 20 2>&1 | grep -E "step|force " | sed "s/^/[$o] $pot /"
  done
done 2>&1 | tee gpurun_out/r2a_ab.txt
for o in "nl_n3=0" ""; do
  COMD_OPTS=$o timeout 900 python bench.py --workload eam_strong --steps 20 --warmup 3 --no-cpu-baseline --e2e-steps 1 2>&1 | grep '^{' | python -c "
import sys,json
for l in sys.stdin:
    j=json.loads(l); print('[$o] eam_strong ms/step %.3f value %.4g force_ms %.3f frac %.4f'%(j['ms_per_step'],j['value'],j['roofline']['force_ms_per_step'],j['roofline']['frac']))"
done 2>&1 | tee gpurun_out/r2a_bench.txt
