#!/bin/bash
# A/B of the halo/compute overlap at N ranks: same workload with overlap off/on, phase timing, energies must agree bit for bit
set -u
N=${1:-2}; TAG=${2:-ovl}; mkdir -p gpurun_out
for wl in ${WLS:-eam_strong}; do
  for o in ${OPTS:-overlap=0,phase_timing=1 overlap=1,phase_timing=1 overlap=0 overlap=1}; do
    COMD_OPTS=$o COMD_NL=1 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
       bench.py --gpus $N --steps ${STEPS:-20} --warmup 3 --workload $wl --no-cpu-baseline --e2e-steps 1 > gpurun_out/${TAG}_n${N}_${wl}_${o//[=,]/_}.log 2>&1
    echo "$wl [$o] N=$N rc=$?"
    grep -A30 "comd phase timing" gpurun_out/${TAG}_n${N}_${wl}_${o//[=,]/_}.log | head -32
    grep "^{" gpurun_out/${TAG}_n${N}_${wl}_${o//[=,]/_}.log | python -c "
import sys,json
for l in sys.stdin:
    j=json.loads(l); print('  ms/step %.3f value %.4g force_ms %.3f launches %d E0 %.12f E1 %.12f'%(j['ms_per_step'],j['value'],j['roofline']['force_ms_per_step'],j['gpu_launches'],j['check']['e_total_per_atom_before'],j['check']['e_total_per_atom_after']))
" || tail -5 gpurun_out/${TAG}_n${N}_${wl}_${o//[=,]/_}.log
    grep -q "^{" gpurun_out/${TAG}_n${N}_${wl}_${o//[=,]/_}.log || tail -15 gpurun_out/${TAG}_n${N}_${wl}_${o//[=,]/_}.log
  done
done
