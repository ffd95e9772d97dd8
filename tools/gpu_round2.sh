#!/bin/bash
# round-end style session on 1 GPU: tests, smoke, default bench (+ reference arm), ncu launch list of the bench command
set -u
mkdir -p gpurun_out
TAG=${1:-r}
python -m pytest tests -m gpu -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/${TAG}_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
python bench.py > gpurun_out/${TAG}_bench_default.log 2>&1; echo "bench rc=$?"; tail -c 2500 gpurun_out/${TAG}_bench_default.log
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${TAG}_bench_reference.log 2>&1; echo "ref rc=$?"; tail -c 900 gpurun_out/${TAG}_bench_reference.log
python bench.py --neighbor-list 0 --no-cpu-baseline > gpurun_out/${TAG}_bench_linkcell.log 2>&1; echo "linkcell rc=$?"
python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/${TAG}_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${TAG}_bench_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/${TAG}_ncu.log 2>&1
echo "ncu rc=$?"
