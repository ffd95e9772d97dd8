#!/bin/bash
# round 2: N GPUs ($1): A/B of the overlap scheme and the sync-free step on C5, then the driver's default invocation
set -u
N=${1:-2}
mkdir -p gpurun_out
T=r2k
run() { timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 bench.py --gpus $N "${@:2}"; }
port=29520
IFS=";" read -ra VARIANTS <<< "${AB:-nl_speculate=0,overlap_mode=0;nl_speculate=0;overlap_mode=0;}"
AB=${AB:-;}; [ "$AB" != "${AB%;}" ] && VARIANTS+=("")
for o in "${VARIANTS[@]}"; do
  port=$((port+1))
  COMD_OPTS=$o run $port --no-cpu-baseline --no-verify --no-secondary --e2e-steps 1 2>/dev/null | grep '^{' > gpurun_out/${T}_ab_n${N}.json
  python - <<P
import json
j=json.load(open('gpurun_out/${T}_ab_n${N}.json'))
print('[N=$N $o] ms/step %.3f force %.3f (every-step window %.3f ms/step) clocks %s'%(j['ms_per_step'], j['roofline']['force_ms_per_step'], 1e3*j['config']['atoms']/j['value_energies_every_step'], j['clocks']))
print('   wall per 10-step call', j['config']['list_in_timed_window']['host_wall_ms_per_call'])
P
done 2>&1 | tee gpurun_out/${T}_ab_n${N}.txt
( time run 29540 ) > gpurun_out/${T}_bench_n$N.log 2>&1
grep '^{' gpurun_out/${T}_bench_n$N.log > gpurun_out/${T}_bench_n$N.json
python - <<P
import json
d=json.load(open('gpurun_out/${T}_bench_n$N.json'))
print('N=$N default: ms/step', d['ms_per_step'], 'value %.4g'%d['value'], 'frac', d['roofline']['frac'], 'e2e %.4g'%d['e2e']['value'])
print(json.dumps(d['check']['parity_vs_oracle'])[:2500])
print({k:(v['ms_per_step'],v['force_ms_per_step']) for k,v in d['secondary'].items()})
P
