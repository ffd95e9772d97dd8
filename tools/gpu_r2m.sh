#!/bin/bash
# round 2: N GPUs ($1): force-kernel variants on C5 at the driver's K/W (20 / 5) and at 100 steps, per-phase medians
set -u
N=${1:-8}
mkdir -p gpurun_out
T=r2m
run() { timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 bench.py --gpus $N "${@:2}"; }
port=29600
IFS=";" read -ra VARIANTS <<< "${AB:-;nl_block=40402;phase_timing=1}"
for o in "${VARIANTS[@]}"; do
  port=$((port+1))
  COMD_OPTS=$o run $port --no-cpu-baseline --no-verify --no-secondary --e2e-steps 1 ${BENCH_ARGS:-} 2> gpurun_out/${T}_err_n${N}.txt | grep '^{' > gpurun_out/${T}_ab_n${N}.json
  grep -A3 "comd phase timing" gpurun_out/${T}_err_n${N}.txt | head -8 | cut -c1-400
  python - <<P
import json
j=json.load(open('gpurun_out/${T}_ab_n${N}.json'))
print('[N=$N $o] steps %d ms/step %.3f force %.3f (every-step window %.3f ms/step) list %s'%(j['steps'], j['ms_per_step'], j['roofline']['force_ms_per_step'], 1e3*j['config']['atoms']/j['value_energies_every_step'], {k:v for k,v in j['config']['list_in_timed_window'].items()}))
P
done 2>&1 | tee gpurun_out/${T}_ab_n${N}.txt
