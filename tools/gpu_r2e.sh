#!/bin/bash
# round 2: GPU tests, then the driver's own bench invocations (default line with every leg, reference arm)
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -12 | tee gpurun_out/r2e_pytest.txt
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -5 | tee gpurun_out/r2e_smoke.txt
( time timeout 1500 python bench.py ) > gpurun_out/r2e_bench_n1.log 2>&1
grep '^{' gpurun_out/r2e_bench_n1.log > gpurun_out/r2e_bench_n1.json
tail -4 gpurun_out/r2e_bench_n1.log | cut -c1-600
( time timeout 600 python bench.py --impl reference --steps 3 --warmup 1 ) > gpurun_out/r2e_ref_n1.log 2>&1
tail -5 gpurun_out/r2e_ref_n1.log | cut -c1-600
