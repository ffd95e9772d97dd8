#!/bin/bash
# GPU session: parity tests, 864k-atom force timings, C5 + C4 bench lines.  usage: gpu_v4.sh TAG [sweep]
set -u
mkdir -p gpurun_out
TAG=${1:-v4}
python -m pytest tests -m gpu -q -x > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/${TAG}_pytest.log
export COMD_NL=1
for o in "" ${2:-}; do
  COMD_OPTS=$o timeout 300 python tools/prof_force.py setfl 60 5 20 2>&1 | grep -E "force|step" | sed "s/^/[$o] /"
done
COMD_OPTS="" timeout 300 python tools/prof_force.py lj 60 5 20 2>&1 | grep -E "force|step"
for wl in eam_strong lj_weak; do
  timeout 900 python bench.py --workload $wl --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_bench_${wl}.log 2>&1; echo "bench $wl rc=$?"
  grep '^{' gpurun_out/${TAG}_bench_${wl}.log | python -c "
import sys,json
for l in sys.stdin:
    j=json.loads(l); print('$wl ms/step %.3f value %.4g force_ms %.3f frac %.4f e2e %.4g launches %d'%(j['ms_per_step'],j['value'],j['roofline']['force_ms_per_step'],j['roofline']['frac'],j['e2e']['value'],j['gpu_launches']))"
done
