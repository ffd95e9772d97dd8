#!/bin/bash
# One GPU session: parity tests, then ncu (launch list + full capture of the force kernels) on a small case.
set -u
mkdir -p gpurun_out
TAG=${1:-prof}; POT=${2:-setfl}; N=${3:-60}; PAT=${4:-k_force_tile}
python -m pytest tests -m gpu -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/${TAG}_pytest.log
python tools/prof_force.py $POT $N 3 3 > gpurun_out/${TAG}_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/${TAG}_launches.csv python tools/prof_force.py $POT $N 3 3 > gpurun_out/${TAG}_ncu1.log 2>&1
cat gpurun_out/${TAG}_plain.log
python tools/prof_force.py $POT $N 3 3 > gpurun_out/${TAG}_plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:$PAT -s 4 -c 2 -o gpurun_out/${TAG}_full python tools/prof_force.py $POT $N 3 3 > gpurun_out/${TAG}_ncu2.log 2>&1
tail -3 gpurun_out/${TAG}_ncu2.log
