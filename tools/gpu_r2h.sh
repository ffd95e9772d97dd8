#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 600 python tools/f32_probe.py f32 2>&1 | tail -15 | tee gpurun_out/r2h_probe_f32.txt
python -m pytest tests/test_gpu_precision.py -q -m gpu > gpurun_out/r2h_precision.txt 2>&1
grep -E "^E  |^(FAILED|ERROR)|passed|failed" gpurun_out/r2h_precision.txt | cut -c1-260 | head -70
