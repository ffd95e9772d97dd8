#!/bin/bash
# round 2: re-validation after the block-order tables took over the interior / boundary enumeration
set -u
mkdir -p gpurun_out
T=r2t
python -m pytest tests -m gpu -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${T}_pytest.log | cut -c1-200
timeout 900 python bench.py --steps 100 --no-cpu-baseline --no-verify --no-secondary --e2e-steps 1 2>/dev/null | grep '^{' > gpurun_out/${T}_bench_k100.json
python - <<P
import json
j=json.load(open('gpurun_out/${T}_bench_k100.json'))
print('k100 ms/step %.3f value %.4g force %.3f frac %.4f'%(j['ms_per_step'], j['value'], j['roofline']['force_ms_per_step'], j['roofline']['frac']))
P
