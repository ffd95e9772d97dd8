#!/bin/bash
# multi-GPU session: bench.py under torchrun at N ranks (as the driver launches it)
set -u
N=${1:-2}; TAG=${2:-multi}
mkdir -p gpurun_out
for wl in ${WLS:-eam_60 eam_strong lj_weak}; do
  for nl in ${NLS:-0 1}; do
    COMD_NL=$nl timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
       bench.py --gpus $N --steps ${STEPS:-20} --warmup 3 --workload $wl --no-cpu-baseline > gpurun_out/${TAG}_n${N}_${wl}_nl${nl}.log 2>&1
    echo "$wl nl=$nl N=$N rc=$?"
    grep "^{" gpurun_out/${TAG}_n${N}_${wl}_nl${nl}.log | python -c "
import sys,json
for l in sys.stdin:
    j=json.loads(l); print('  ms/step %.3f value %.4g e2e %.4g launches %d E0 %.9f E1 %.9f'%(j['ms_per_step'],j['value'],j['e2e']['value'],j['gpu_launches'],j['check']['e_total_per_atom_before'],j['check']['e_total_per_atom_after']))
" || tail -5 gpurun_out/${TAG}_n${N}_${wl}_nl${nl}.log
    grep -q "^{" gpurun_out/${TAG}_n${N}_${wl}_nl${nl}.log || tail -15 gpurun_out/${TAG}_n${N}_${wl}_nl${nl}.log
  done
done
