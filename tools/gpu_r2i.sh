#!/bin/bash
# round 2, re-entry session: GPU tests, smoke, the driver's default bench line + reference arm, the ncu launch list of a
# short default bench, and one ncu --set full capture of the once-per-pair list kernels (EAM setfl 60^3)
set -u
mkdir -p gpurun_out
T=r2i
python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/${T}_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2 | tee gpurun_out/${T}_smoke.txt
( time timeout 1500 python bench.py ) > gpurun_out/${T}_bench_n1.log 2>&1; echo "bench rc=$?"
grep '^{' gpurun_out/${T}_bench_n1.log > gpurun_out/${T}_bench_n1.json
tail -4 gpurun_out/${T}_bench_n1.log | cut -c1-1500
( time timeout 600 python bench.py --impl reference --steps 3 --warmup 1 ) > gpurun_out/${T}_ref_n1.log 2>&1
grep '^{' gpurun_out/${T}_ref_n1.log > gpurun_out/${T}_ref_n1.json; tail -4 gpurun_out/${T}_ref_n1.log | cut -c1-600
# launch list of the timed region's command (short), after it exited 0 without ncu
timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-verify --no-secondary --e2e-steps 1 > gpurun_out/${T}_plain.log 2>&1 && \
timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/${T}_launches_c5.csv python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-verify --no-secondary --e2e-steps 1 > gpurun_out/${T}_ncu_ll.log 2>&1
echo "ncu launch list rc=$?"; python tools/launch_summary.py gpurun_out/${T}_launches_c5.csv | tee gpurun_out/${T}_launches_c5.txt | head -40
export COMD_NL=1
timeout 300 python tools/prof_force.py setfl 60 5 20 2>&1 | tail -4
ncu --set full --clock-control none --import-source on -k regex:'k_force_n3|k_nl_build' -s 3 -c 6 -o gpurun_out/${T}_n3_full python tools/prof_force.py setfl 60 3 12 > gpurun_out/${T}_ncu.log 2>&1
tail -3 gpurun_out/${T}_ncu.log
ncu -i gpurun_out/${T}_n3_full.ncu-rep --page raw --csv > gpurun_out/${T}_n3_raw.csv 2>/dev/null
python tools/ncu_summary.py gpurun_out/${T}_n3_raw.csv | tee gpurun_out/${T}_n3_summary.txt | head -80
