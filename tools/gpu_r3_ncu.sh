#!/bin/bash
# Round 3: ncu --set full of the kernels round 2 changed but did not capture (missing #4):
# k_force_n3<EAMFE> beside <RHO>/<EAMF>, k_kick_drift, and the -DCOMD_SINGLE twins; EAM Cu01 setfl 60^3 (864 k atoms), list path.
# Each capture runs only after the same command has exited 0 without ncu.  usage: bash tools/gpu_r3_ncu.sh
set -u
mkdir -p gpurun_out
export COMD_NL=1
for prec in f64 f32; do
  export COMD_PREC=$prec
  TAG=r3_n3_${prec}
  python tools/prof_force.py setfl 60 3 12 > gpurun_out/${TAG}_plain.log 2>&1 || { cat gpurun_out/${TAG}_plain.log; exit 1; }
  cat gpurun_out/${TAG}_plain.log
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_force_n3|k_kick_drift' -s 2 -c 12 \
      -o gpurun_out/${TAG}_full -f python tools/prof_force.py setfl 60 3 12 > gpurun_out/${TAG}_ncu.log 2>&1
  tail -2 gpurun_out/${TAG}_ncu.log
  ncu -i gpurun_out/${TAG}_full.ncu-rep --page raw --csv > gpurun_out/${TAG}_raw.csv 2>/dev/null
  python tools/ncu_summary.py gpurun_out/${TAG}_raw.csv > gpurun_out/${TAG}_summary.txt
  ls -la gpurun_out/${TAG}_full.ncu-rep
  [ $(stat -c %s gpurun_out/${TAG}_full.ncu-rep) -gt 25000000 ] && rm -f gpurun_out/${TAG}_full.ncu-rep   # keep the merge under its size cap
done
# launch list of the default bench command (per-kernel share of the step), after a plain run of the same command
python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-secondary > gpurun_out/r3_launch_plain.json 2> gpurun_out/r3_launch_plain.err || exit 1
timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r3_launches_c5.csv \
    python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-secondary > gpurun_out/r3_launch_ncu.log 2>&1
python tools/launch_summary.py gpurun_out/r3_launches_c5.csv > gpurun_out/r3_launches_c5.txt 2>&1
tail -25 gpurun_out/r3_launches_c5.txt
