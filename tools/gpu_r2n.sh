#!/bin/bash
# round 2: GPU suite (Hilbert order, full-size tests, build sweep without the skipped candidates), list-build timing, one ncu capture of the build
set -u
mkdir -p gpurun_out
T=r2n
python -m pytest tests -m gpu -q -x > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?"
grep -E "^E  |^(FAILED|ERROR)|passed|failed" gpurun_out/${T}_pytest.log | cut -c1-300 | head -40
export COMD_NL=1
timeout 600 python tools/prof_build.py setfl 160 2>&1 | tail -3 | tee gpurun_out/${T}_build_c5.txt
timeout 600 python tools/prof_build.py setfl 60 2>&1 | tail -3 | tee -a gpurun_out/${T}_build_c5.txt
for o in "" "nl_hilbert=1"; do COMD_OPTS=$o timeout 300 python tools/prof_force.py setfl 160 3 20 2>&1 | grep -E "step|force " | sed "s/^/[$o] /"; done | tee gpurun_out/${T}_hilbert_c5.txt
ncu --set full --clock-control none --import-source on -k regex:'k_nl_build' -c 1 -o gpurun_out/${T}_build_full python tools/prof_force.py setfl 60 1 0 > gpurun_out/${T}_ncu.log 2>&1
tail -2 gpurun_out/${T}_ncu.log
ncu -i gpurun_out/${T}_build_full.ncu-rep --page raw --csv > gpurun_out/${T}_build_raw.csv 2>/dev/null
python tools/ncu_summary.py gpurun_out/${T}_build_raw.csv | tee gpurun_out/${T}_build_summary.txt | head -30
