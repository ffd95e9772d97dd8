#!/bin/bash
# ncu --set full capture of the list kernels (and the list build) on EAM setfl 60^3; usage: gpu_ncu.sh TAG [opts]
set -u
mkdir -p gpurun_out
TAG=${1:-ncu}; export COMD_OPTS=${2:-}; export COMD_NL=1
python tools/prof_force.py setfl 60 3 12 > gpurun_out/${TAG}_plain.log 2>&1 || { cat gpurun_out/${TAG}_plain.log; exit 1; }
cat gpurun_out/${TAG}_plain.log
ncu --set full --clock-control none --import-source on -k regex:'k_force_nl|k_nl_build' -s 3 -c 6 -o gpurun_out/${TAG}_full python tools/prof_force.py setfl 60 3 12 > gpurun_out/${TAG}_ncu.log 2>&1
tail -3 gpurun_out/${TAG}_ncu.log
