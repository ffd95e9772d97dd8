#!/bin/bash
# round 2: the driver's multi-GPU bench invocation on N GPUs (N = $1), both arms
set -u
N=${1:-2}
mkdir -p gpurun_out
( time timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N ) > gpurun_out/r2f_bench_n$N.log 2>&1
grep '^{' gpurun_out/r2f_bench_n$N.log > gpurun_out/r2f_bench_n$N.json
tail -12 gpurun_out/r2f_bench_n$N.log | cut -c1-400
python - <<P
import json
d=json.load(open('gpurun_out/r2f_bench_n$N.json'))
print('N=$N ms/step', d['ms_per_step'], 'value %.4g'%d['value'], 'frac', d['roofline']['frac'])
print(json.dumps(d['check']['parity_vs_oracle'])[:2500])
P
( time timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus $N --steps 3 --warmup 1 ) > gpurun_out/r2f_ref_n$N.log 2>&1
grep '^{' gpurun_out/r2f_ref_n$N.log | cut -c1-300
