#!/bin/bash
# The measurement matrix of BASELINE.md at N GPUs of one box: every config, native arm and the
# torch+NCCL arm in the same run (bench.py times both), value tests of the data plane at this
# world size, and the Hogwild step breakdown.  Usage (through gpurun --gpus N): bash tools/matrix.sh N
set -u
N=${1:-1}
O=gpurun_out/matrix_n$N
mkdir -p $O
if [ "$N" -gt 1 ]; then
  RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
else
  RUN="python"
fi
port=29700
bench() {   # name, args...
  name=$1; shift
  port=$((port + 1))
  if [ "$N" -gt 1 ]; then P="--master-port $port"; else P=""; fi
  timeout 900 $RUN $P bench.py --gpus $N "$@" 2> $O/$name.err | tail -1 > $O/$name.json
  python - "$O/$name.json" "$name" <<'PY'
import json, sys
try:
    d = json.load(open(sys.argv[1]))
    b = d.get("baseline") or {}
    print("%-22s N=%d native %.5g (%.3f ms) e2e %.5g | torch %.5g (%.3f ms) | ratio %.3f e2e-ratio %s | loss %.4f / %.4f %s" % (
        sys.argv[2], d["n_gpus"], d["value"], d["ms_per_step"], (d.get("e2e") or {}).get("value", 0),
        b.get("value", 0), b.get("ms_per_step", 0), d.get("vs_baseline") or 0,
        ("%.3f" % b["e2e_ratio"]) if b.get("e2e_ratio") else "-", d.get("final_loss") or 0,
        b.get("final_loss") or 0, (d["config"].get("nvlink") or "")))
except Exception as e:
    print("%-22s FAILED (%s)" % (sys.argv[2], e))
PY
}
if [ "$N" -gt 1 ]; then
  for nv in 0 1; do
    port=$((port + 1))
    DTG_NVLS=$nv timeout 400 $RUN --master-port $port tests/gpu_value_worker.py > $O/value_nvls$nv.log 2>&1
    echo "== value tests, world $N, NVLS=$nv (rc=$?)"
    grep RESULT $O/value_nvls$nv.log | python -c "
import sys, json
for l in sys.stdin:
    r = json.loads(l[7:])
    for k, v in r.items(): print('  %-42s %s' % (k, 'ok' if v.get('ok') else 'FAIL ' + json.dumps(v)[:500]))
"
  done
  port=$((port + 1)); timeout 300 $RUN --master-port $port tools/step_breakdown.py 2>&1 | grep BREAKDOWN | tee $O/breakdown_striped.json | cut -c1-900
  port=$((port + 1)); timeout 300 $RUN --master-port $port tools/step_breakdown.py --ps_layout gpu0 2>&1 | grep BREAKDOWN | tee $O/breakdown_gpu0.json | cut -c1-900
fi
bench hogwild_striped --steps 40 --warmup 5
if [ "$N" -gt 1 ]; then
  [ "$N" -lt 8 ] && bench hogwild_gpu0 --steps 40 --warmup 5 --ps_layout gpu0
  bench hogwild_gpu0_ps_only --steps 40 --warmup 5 --ps_layout gpu0_ps_only
  [ "$N" -lt 8 ] && bench hogwild_gpu0_ps_only_fusedpush --steps 40 --warmup 5 --ps_layout gpu0_ps_only --async_push 0 --no_baseline
fi
bench gpt2_ssgd_lr --model gpt2_medium --scheme ssgd_lr --steps 8 --warmup 3
bench bert_tower_avg --model bert_large --scheme tower_avg --steps 8 --warmup 3
bench resnet_adag --model resnet50 --scheme adag --steps 3 --warmup 2
bench resnet_downpour --model resnet50 --scheme downpour --steps 3 --warmup 2
grep -l "Error\|error" $O/*.err 2>/dev/null | head
du -sh $O
exit 0
