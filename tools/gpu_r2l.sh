#!/bin/bash
# round 2: are the sporadic +45 ms intervals of the multi-rank timed window tied to the clock sampler?  N GPUs ($1), 4 runs each way
set -u
N=${1:-2}
mkdir -p gpurun_out
T=r2l
run() { timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 bench.py --gpus $N "${@:2}"; }
port=29560
for rep in 1 2 3 4; do for smp in 1 ""; do
  port=$((port+1))
  COMD_BENCH_NO_SAMPLER=$smp run $port --no-cpu-baseline --no-verify --no-secondary --e2e-steps 1 2>/dev/null | grep '^{' > gpurun_out/${T}_n${N}.json
  python - <<P
import json
j=json.load(open('gpurun_out/${T}_n${N}.json'))
print('[N=$N no_sampler=$smp] ms/step %.3f force %.3f | wall per 10-step call'%(j['ms_per_step'], j['roofline']['force_ms_per_step']), j['config']['list_in_timed_window']['host_wall_ms_per_call'], '| every-step window', j['config']['list_in_every_step_window']['host_wall_ms_per_call'])
P
done; done 2>&1 | tee gpurun_out/${T}_sampler_n${N}.txt
