#!/bin/bash
# GPU test suite only; the tail of the log comes back on stdout, the full log in gpurun_out/
mkdir -p gpurun_out
python -m pytest tests -m gpu -q ${PYTEST_ARGS:-} > gpurun_out/${1:-t}_pytest.log 2>&1; echo "pytest rc=$?"
grep -E "^E  |^(FAILED|ERROR)|passed|failed" gpurun_out/${1:-t}_pytest.log | cut -c1-300 | head -60
