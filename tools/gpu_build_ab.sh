#!/bin/bash
# parity tests + bench lines after a list-build change: usage gpu_build_ab.sh TAG
set -u
mkdir -p gpurun_out
TAG=${1:-ab}
python -m pytest tests -m gpu -q -x > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${TAG}_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
for wl in eam_strong lj_weak; do
  timeout 600 python bench.py --workload $wl --steps 40 --warmup 3 --no-cpu-baseline --e2e-steps 1 > gpurun_out/${TAG}_bench_${wl}.log 2>&1; echo "$wl rc=$?"
  grep '^{' gpurun_out/${TAG}_bench_${wl}.log | python -c "
import sys,json
for l in sys.stdin:
    j=json.loads(l); print('$wl ms/step %.3f value %.4g force_ms %.3f frac %.4f builds %s'%(j['ms_per_step'],j['value'],j['roofline']['force_ms_per_step'],j['roofline']['frac'],j['config'].get('neighbor_list')))"
done
