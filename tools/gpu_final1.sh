#!/bin/bash
# final single-GPU session of the round: parity tests, smoke, bench lines (C5, C4, C3 both paths), reference arm, ncu full of the list kernels
set -u
mkdir -p gpurun_out
TAG=${1:-fin}
python -m pytest tests -m gpu -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${TAG}_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
python bench.py > gpurun_out/${TAG}_bench_default.log 2>&1; echo "bench default rc=$?"; grep '^{' gpurun_out/${TAG}_bench_default.log | cut -c1-300
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${TAG}_bench_reference.log 2>&1; echo "reference rc=$?"; grep '^{' gpurun_out/${TAG}_bench_reference.log | cut -c1-700
for wl in lj_weak eam_60; do
  python bench.py --workload $wl --no-cpu-baseline > gpurun_out/${TAG}_bench_${wl}.log 2>&1; echo "$wl rc=$?"; grep '^{' gpurun_out/${TAG}_bench_${wl}.log | cut -c1-220
done
python bench.py --workload eam_60 --neighbor-list 0 --no-cpu-baseline > gpurun_out/${TAG}_bench_eam_60_linkcell.log 2>&1; echo "eam_60 link-cell rc=$?"; grep '^{' gpurun_out/${TAG}_bench_eam_60_linkcell.log | cut -c1-220
bash tools/gpu_ncu.sh ${TAG}
