#!/bin/bash
# quick A/B of list-kernel options on 864k atoms (+ optional bench): usage gpu_quick.sh "opt1 opt2 ..." [bench workloads]
set -u
export COMD_NL=1
for o in "" $1; do
  for pot in setfl lj; do
    COMD_OPTS=$o timeout 300 python tools/prof_force.py $pot 60 5 20 2>&1 | grep -E "step" | sed "s/^/[$o] $pot /"
  done
done
for wl in ${2:-}; do
  for o in "" $1; do
  COMD_OPTS=$o timeout 900 python bench.py --workload $wl --steps 20 --warmup 3 --no-cpu-baseline --e2e-steps 1 2>&1 | grep '^{' | python -c "
import sys,json
for l in sys.stdin:
    j=json.loads(l); print('[$o] $wl ms/step %.3f value %.4g force_ms %.3f frac %.4f'%(j['ms_per_step'],j['value'],j['roofline']['force_ms_per_step'],j['roofline']['frac']))"
  done
done
