#!/bin/bash
# ncu launch list (gpu__time_duration) of the default bench command, after the same command exited 0 without ncu
set -u
mkdir -p gpurun_out
TAG=${1:-ll}; WL=${2:-eam_strong}
timeout 900 python bench.py --workload $WL --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_bench_${WL}.log 2>&1 || { tail -5 gpurun_out/${TAG}_bench_${WL}.log; exit 1; }
grep '^{' gpurun_out/${TAG}_bench_${WL}.log | cut -c1-200
timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/${TAG}_launches_${WL}.csv python bench.py --workload $WL --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_ncu_${WL}.log 2>&1
echo "ncu rc=$?"; python tools/launch_summary.py gpurun_out/${TAG}_launches_${WL}.csv
