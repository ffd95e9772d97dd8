#!/bin/bash
# round 2: launch-shape sweep of the once-per-pair kernels on C5 (one GPU, 40 steps each)
set -u
mkdir -p gpurun_out
T=r2u
for o in "" "nl_chunks_per_cta=8" "nl_chunks_per_cta=32" "nl_threads=384" "nl_threads=416" "nl_block=40404,nl_chunks_per_cta=4"; do
  COMD_OPTS=$o timeout 600 python bench.py --steps 40 --warmup 5 --no-cpu-baseline --no-verify --no-secondary --e2e-steps 1 2>/dev/null | grep '^{' > gpurun_out/${T}_ab.json
  python - <<P
import json
j=json.load(open('gpurun_out/${T}_ab.json'))
print('[$o] ms/step %.3f force %.3f (every-step window: force %.3f) builds %d'%(j['ms_per_step'], j['roofline']['force_ms_per_step'], j['roofline']['force_ms_per_step_energies_every_step'], j['config']['list_in_timed_window']['builds']))
P
done 2>&1 | tee gpurun_out/${T}_launch_shape_c5.txt
