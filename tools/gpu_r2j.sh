#!/bin/bash
# round 2: GPU tests after the sync-free step / one-shot interior launches / f32 host build, then a short default bench
set -u
mkdir -p gpurun_out
T=r2j
python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/${T}_pytest.log | cut -c1-300
for o in "nl_speculate=0" ""; do
  COMD_OPTS=$o timeout 900 python bench.py --no-cpu-baseline --no-secondary --e2e-steps 1 2>&1 | grep '^{' > gpurun_out/${T}_bench_"${o:-default}".json
  python - <<P
import json
j=json.load(open('gpurun_out/${T}_bench_${o:-default}.json'))
print('[$o] ms/step %.3f force %.3f frac %.4f list %s parity %s clocks %s'%(j['ms_per_step'], j['roofline']['force_ms_per_step'], j['roofline']['frac'], {k:v for k,v in j['config']['list_in_timed_window'].items() if k!='host_wall_ms_per_call'}, j['check']['parity_vs_oracle']['pass'], j['clocks']))
print(j['config']['list_in_timed_window']['host_wall_ms_per_call'])
P
done
