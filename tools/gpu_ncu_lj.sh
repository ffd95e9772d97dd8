#!/bin/bash
# ncu --set full of the LJ list kernel and the list build (864k atoms)
set -u
mkdir -p gpurun_out
TAG=${1:-lj}; export COMD_NL=1
python tools/prof_force.py lj 60 3 12 > gpurun_out/${TAG}_plain.log 2>&1 || { cat gpurun_out/${TAG}_plain.log; exit 1; }
cat gpurun_out/${TAG}_plain.log
ncu --set full --clock-control none --import-source on -k regex:'k_force_nl|k_nl_build' -c 3 -o gpurun_out/${TAG}_full python tools/prof_force.py lj 60 3 12 > gpurun_out/${TAG}_ncu.log 2>&1
tail -2 gpurun_out/${TAG}_ncu.log
