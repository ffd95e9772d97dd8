#!/bin/bash
# round 2: GPU tests, timings of the once-per-pair kernels, then one ncu --set full capture of them (EAM setfl 60^3)
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -12 | tee gpurun_out/r2d_pytest.txt
export COMD_NL=1
for o in "" "nl_lanes=8" "nl_n3=0"; do
  for pot in setfl lj; do
    COMD_OPTS=$o timeout 300 python tools/prof_force.py $pot 60 5 20 2>&1 | grep -E "step|force " | sed "s/^/[$o] $pot /"
  done
done 2>&1 | tee gpurun_out/r2d_ab.txt
COMD_OPTS="" timeout 900 python bench.py --workload eam_strong --steps 20 --warmup 3 --no-cpu-baseline --no-verify --no-secondary --e2e-steps 1 2>&1 | grep '^{' > gpurun_out/r2d_bench_c5.json
python -c "
import json; j=json.load(open('gpurun_out/r2d_bench_c5.json')); print('eam_strong ms/step %.3f value %.4g force_ms %.3f frac %.4f'%(j['ms_per_step'],j['value'],j['roofline']['force_ms_per_step'],j['roofline']['frac']))"
ncu --set full --clock-control none --import-source on -k regex:'k_force_n3|k_nl_build' -s 3 -c 6 -o gpurun_out/r2d_n3_full python tools/prof_force.py setfl 60 3 12 > gpurun_out/r2d_ncu.log 2>&1
tail -3 gpurun_out/r2d_ncu.log
ncu -i gpurun_out/r2d_n3_full.ncu-rep --page raw --csv > gpurun_out/r2d_n3_raw.csv 2>/dev/null
python tools/ncu_summary.py gpurun_out/r2d_n3_raw.csv | tee gpurun_out/r2d_n3_summary.txt
