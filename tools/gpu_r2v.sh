#!/bin/bash
# round 2: list build after the instruction diet (no per-row divisions, range-compare classification): tests, timing
set -u
mkdir -p gpurun_out
T=r2v
python -m pytest tests -m gpu -q -x -k "neighbor_list or precision or smoke or c3_full or c4_c5" > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${T}_pytest.log | cut -c1-200
export COMD_NL=1
timeout 600 python tools/prof_build.py setfl 160 2>&1 | tail -1 | tee gpurun_out/${T}_build.txt
timeout 600 python tools/prof_build.py setfl 60 2>&1 | tail -1 | tee -a gpurun_out/${T}_build.txt
timeout 600 python tools/prof_build.py lj 100 2>&1 | tail -1 | tee -a gpurun_out/${T}_build.txt
