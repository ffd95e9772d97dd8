#!/bin/bash
# round 2: one GPU: block shape and lanes-per-row A/B on C5 and C3 (100 steps), shared-atomics microbenchmark
set -u
mkdir -p gpurun_out
T=r2o
for wl in eam_strong eam_60; do
for o in "" "nl_block=40402" "nl_lanes=8"; do
  COMD_OPTS=$o timeout 600 python bench.py --workload $wl --no-cpu-baseline --no-verify --no-secondary --e2e-steps 1 2>/dev/null | grep '^{' > gpurun_out/${T}_ab.json
  python - <<P
import json
j=json.load(open('gpurun_out/${T}_ab.json'))
print('[$wl $o] steps %d ms/step %.3f force %.3f (every-step window %.3f ms/step, force %.3f) builds %d'%(j['steps'], j['ms_per_step'], j['roofline']['force_ms_per_step'], 1e3*j['config']['atoms']/j['value_energies_every_step'], j['roofline']['force_ms_per_step_energies_every_step'], j['config']['list_in_timed_window']['builds']))
P
done; done 2>&1 | tee gpurun_out/${T}_block_lanes_ab.txt
nvcc -O3 -gencode arch=compute_100a,code=sm_100a tools/micro/atoms_bench.cu -o gpurun_out/atoms_bench && ./gpurun_out/atoms_bench | tee gpurun_out/${T}_micro_atoms.txt; rm -f gpurun_out/atoms_bench
