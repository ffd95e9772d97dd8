#!/bin/bash
# round 2, final single-GPU pass: GPU suite, smoke, the driver's own invocations (K = 20, W = 5) of both arms, a 100-step line,
# the ncu launch list of the bench command and one ncu --set full capture of the once-per-pair kernels of the final code
set -u
mkdir -p gpurun_out
T=r2s
python -m pytest tests -m gpu -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${T}_pytest.log | cut -c1-200
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2 | tee gpurun_out/${T}_smoke.txt
( time timeout 1500 python bench.py --gpus 1 --steps 20 --warmup 5 ) > gpurun_out/${T}_bench_k20.log 2>&1; echo "bench rc=$?"
grep '^{' gpurun_out/${T}_bench_k20.log > gpurun_out/${T}_bench_k20.json; tail -4 gpurun_out/${T}_bench_k20.log | cut -c1-400
( time timeout 600 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 ) > gpurun_out/${T}_ref.log 2>&1
grep '^{' gpurun_out/${T}_ref.log > gpurun_out/${T}_ref.json; tail -4 gpurun_out/${T}_ref.log | cut -c1-300
timeout 900 python bench.py --steps 100 --no-cpu-baseline --no-verify 2>/dev/null | grep '^{' > gpurun_out/${T}_bench_k100.json
python - <<P
import json
for f in ('k20','k100'):
    j=json.load(open('gpurun_out/${T}_bench_%s.json'%f))
    print(f, 'ms/step %.3f value %.4g force %.3f frac %.4f (every step: %.4f) e2e %.4g list %s'%(j['ms_per_step'], j['value'], j['roofline']['force_ms_per_step'], j['roofline']['frac'], j['roofline']['frac_energies_every_step'], j['e2e']['value'], {k:v for k,v in j['config']['list_in_timed_window'].items() if k!='host_wall_ms_per_call'}))
    print('   secondary', {k:(round(v['ms_per_step'],3),round(v['force_ms_per_step'],3)) for k,v in (j['secondary'] or {}).items()})
P
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-verify --no-secondary --e2e-steps 1 > gpurun_out/${T}_plain.log 2>&1 && \
timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/${T}_launches_c5.csv python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-verify --no-secondary --e2e-steps 1 > gpurun_out/${T}_ncu_ll.log 2>&1
echo "ncu launch list rc=$?"; python tools/launch_summary.py gpurun_out/${T}_launches_c5.csv | tee gpurun_out/${T}_launches_c5.txt | head -14
export COMD_NL=1
ncu --set full --clock-control none --import-source on -k regex:'k_force_n3' -s 3 -c 4 -o gpurun_out/${T}_n3_full python tools/prof_force.py setfl 60 3 12 > gpurun_out/${T}_ncu.log 2>&1
tail -2 gpurun_out/${T}_ncu.log
ncu -i gpurun_out/${T}_n3_full.ncu-rep --page raw --csv > gpurun_out/${T}_n3_raw.csv 2>/dev/null
python tools/ncu_summary.py gpurun_out/${T}_n3_raw.csv > gpurun_out/${T}_n3_summary.txt; head -12 gpurun_out/${T}_n3_summary.txt | cut -c1-160
