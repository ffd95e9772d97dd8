#!/bin/bash
# round 2: precision tests on the GPU (both libraries), then the whole GPU suite
set -u
mkdir -p gpurun_out
python -m pytest tests/test_gpu_precision.py -q -m gpu 2>&1 | tail -40 | tee gpurun_out/r2g_precision.txt
python -m pytest tests -m gpu -x -q 2>&1 | tail -8 | tee gpurun_out/r2g_pytest.txt
