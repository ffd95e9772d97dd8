#!/bin/bash
# round 2: CUDA-graph step: equality tests, then the whole GPU suite, then launch-bound sizes with and without the graph
set -u
mkdir -p gpurun_out
T=r2p
python -m pytest tests -m gpu -q -x -k "graph or sync_free" > gpurun_out/${T}_pytest_graph.log 2>&1; echo "graph tests rc=$?"
grep -E "^E  |^(FAILED|ERROR)|passed|failed" gpurun_out/${T}_pytest_graph.log | cut -c1-300 | head -30
python -m pytest tests -m gpu -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?"
grep -E "^E  |^(FAILED|ERROR)|passed|failed" gpurun_out/${T}_pytest.log | cut -c1-300 | head -30
export COMD_NL=1
for sz in "funcfl 20" "lj 20" "setfl 30" "setfl 60"; do for o in "graph_step=0" "graph_step=1"; do
  COMD_OPTS=$o timeout 300 python tools/prof_force.py $sz 3 400 2>&1 | grep -E "^step" | sed "s/^/[$sz $o] /"
done; done | tee gpurun_out/${T}_graph_ab.txt
