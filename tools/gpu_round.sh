#!/bin/bash
# One GPU session: parity tests, then bench lines for both force variants.  Run under gpurun.
set -u
mkdir -p gpurun_out
TAG=${1:-run}
python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/${TAG}_pytest.log
tail -3 gpurun_out/${TAG}_pytest.log
for wl in eam_60 eam_strong lj_weak; do
  for v in 1 0; do
    COMD_FORCE_VARIANT=$v timeout 600 python bench.py --workload $wl --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_bench_${wl}_v$v.log 2>&1
    echo "$wl v$v rc=$?"; python - <<PY
import json
try:
    l=[x for x in open("gpurun_out/${TAG}_bench_${wl}_v$v.log") if x.startswith("{")][-1]
    j=json.loads(l); print("  ms/step %.3f  value %.4g  force_ms %.3f frac %.4f e2e %.4g"%(j["ms_per_step"],j["value"],j["roofline"]["force_ms_per_step"],j["roofline"]["frac"],j["e2e"]["value"]))
except Exception as ex: print("  no json:",ex)
PY
  done
done
