#!/bin/bash
# GPU session: block-mapped list kernels (v4) -- parity tests, block/threads sweep on 864k atoms, C5 bench.
set -u
mkdir -p gpurun_out
TAG=${1:-v4}
python -m pytest tests -m gpu -q -x > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/${TAG}_pytest.log
export COMD_NL=1
for o in "" "nl_block=40402" "nl_block=40404" "nl_block=40202" "nl_block=80402" "nl_block=20202" "nl_threads=512" "nl_threads=384" "energy_every_step=1"; do
  COMD_OPTS=$o timeout 300 python tools/prof_force.py setfl 60 5 20 2>&1 | grep -E "force|step" | sed "s/^/[$o] /"
done
COMD_OPTS="" timeout 300 python tools/prof_force.py lj 60 5 20 2>&1 | grep -E "force|step"
timeout 900 python bench.py --workload eam_strong --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_bench_c5.log 2>&1; echo "bench rc=$?"
grep '^{' gpurun_out/${TAG}_bench_c5.log | python -c "
import sys,json
for l in sys.stdin:
    j=json.loads(l); print('C5 ms/step %.3f value %.4g force_ms %.3f frac %.4f e2e %.4g'%(j['ms_per_step'],j['value'],j['roofline']['force_ms_per_step'],j['roofline']['frac'],j['e2e']['value']))"
tail -3 gpurun_out/${TAG}_bench_c5.log | cut -c1-600
