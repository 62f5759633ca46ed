#!/bin/bash
# First multi-GPU pass (2 GPUs): value tests, engine tests, step breakdown, bench.  Every
# multi-rank command runs under its own timeout so that a hang costs minutes, not the box.
set -u
N=${1:-2}
O=gpurun_out
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
echo "== value worker, TMA push"; timeout 300 $TR --master-port 29611 tests/gpu_value_worker.py > $O/value_n${N}.log 2>&1; echo "rc=$?"
grep RESULT $O/value_n${N}.log | python -c "
import sys, json
for l in sys.stdin:
    r = json.loads(l[7:])
    for k, v in r.items(): print('  %-42s %s' % (k, 'ok' if v.get('ok') else 'FAIL ' + json.dumps(v)[:600]))
" || tail -30 $O/value_n${N}.log
echo "== value worker, NVLS forced"; DTG_NVLS=1 timeout 300 $TR --master-port 29612 tests/gpu_value_worker.py > $O/value_n${N}_nvls.log 2>&1; echo "rc=$?"
grep RESULT $O/value_n${N}_nvls.log | python -c "
import sys, json
for l in sys.stdin:
    r = json.loads(l[7:])
    for k, v in r.items(): print('  %-42s %s' % (k, 'ok' if v.get('ok') else 'FAIL ' + json.dumps(v)[:600]))
" || tail -30 $O/value_n${N}_nvls.log
echo "== pytest multigpu"; timeout 900 python -m pytest tests/test_gpu_engine.py tests/test_gpu_multi_values.py -m multigpu -q 2>&1 | tail -12
echo "== step breakdown"; timeout 300 $TR --master-port 29613 tools/step_breakdown.py 2>&1 | grep BREAKDOWN | tee $O/breakdown_n${N}.json
echo "== bench"; timeout 600 $TR --master-port 29614 bench.py --gpus $N --steps 30 --warmup 5 2>$O/bench_n${N}.err | tail -1 | tee $O/bench_n${N}.json | cut -c1-1500
tail -3 $O/bench_n${N}.err
exit 0
