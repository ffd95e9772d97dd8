#!/bin/bash
# GPU session for the neighbour-list path: parity tests, bench lines, ncu launch list + full capture.
set -u
mkdir -p gpurun_out
TAG=${1:-nl}
python -m pytest tests -m gpu -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/${TAG}_pytest.log
for wl in eam_strong lj_weak eam_60; do
  COMD_NL=1 timeout 600 python bench.py --workload $wl --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_bench_${wl}_nl.log 2>&1
  echo "$wl nl rc=$?"; python - <<PY
import json
try:
    l=[x for x in open("gpurun_out/${TAG}_bench_${wl}_nl.log") if x.startswith("{")][-1]
    j=json.loads(l); print("  ms/step %.3f  value %.4g  force_ms %.3f frac %.4f e2e %.4g launches %d"%(j["ms_per_step"],j["value"],j["roofline"]["force_ms_per_step"],j["roofline"]["frac"],j["e2e"]["value"],j["gpu_launches"]))
except Exception as ex: print("  no json:",ex); print(open("gpurun_out/${TAG}_bench_${wl}_nl.log").read()[-1500:])
PY
done
export COMD_NL=1
python tools/prof_force.py setfl 60 3 20 > gpurun_out/${TAG}_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/${TAG}_launches.csv python tools/prof_force.py setfl 60 3 20 > gpurun_out/${TAG}_ncu1.log 2>&1
cat gpurun_out/${TAG}_plain.log
python tools/prof_force.py setfl 60 3 0 > gpurun_out/${TAG}_plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_force_nl -s 4 -c 2 -o gpurun_out/${TAG}_full python tools/prof_force.py setfl 60 3 0 > gpurun_out/${TAG}_ncu2.log 2>&1
tail -3 gpurun_out/${TAG}_ncu2.log
