#!/bin/bash
# round 2: N GPUs ($1): parity leg incl. the graph-replayed variant (NCCL send/recv inside the captured step), short timed region
set -u
N=${1:-2}
mkdir -p gpurun_out
T=r2q
( time timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29650 bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline ${BENCH_ARGS:-} ) > gpurun_out/${T}_bench_n$N.log 2>&1
echo "rc=$?"
grep '^{' gpurun_out/${T}_bench_n$N.log > gpurun_out/${T}_bench_n$N.json
python - <<P
import json
d=json.load(open('gpurun_out/${T}_bench_n$N.json'))
print('N=$N K=20 W=5: ms/step', d['ms_per_step'], 'value %.4g'%d['value'], 'frac', d['roofline']['frac'], 'e2e %.4g'%d['e2e']['value'], d['config']['list_in_timed_window'])
p=d['check']['parity_vs_oracle']; print({k:p[k] for k in p if k not in ('config','tolerances')})
print({k:(round(v['ms_per_step'],3),round(v['force_ms_per_step'],3)) for k,v in d['secondary'].items()})
P
tail -3 gpurun_out/${T}_bench_n$N.log | cut -c1-300
