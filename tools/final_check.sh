#!/bin/bash
# One-GPU validation of the tree as it will be judged: GPU test suite, smoke(), one bench line.
# Usage (on a GPU box): bash tools/final_check.sh [extra bench args...]
mkdir -p gpurun_out/final
timeout 400 python -m pytest tests -m gpu -x -q > gpurun_out/final/pytest.txt 2>&1
echo "pytest rc=$?"; tail -n 6 gpurun_out/final/pytest.txt
timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -n 3
if [ $# -gt 0 ]; then
  timeout 400 python bench.py "$@" > gpurun_out/final/bench_extra.json 2> gpurun_out/final/bench_extra.err
  echo "bench rc=$?"; tail -c 1500 gpurun_out/final/bench_extra.json; tail -n 3 gpurun_out/final/bench_extra.err
fi
