N=4
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
DTG_NVLS=1 timeout 300 $RUN --master-port 29801 tests/gpu_value_worker.py 2>&1 | grep RESULT | python -c "
import sys, json
for l in sys.stdin:
    r = json.loads(l[7:]); print('value tests:', sum(1 for v in r.values() if v.get('ok')), 'of', len(r), [k for k,v in r.items() if not v.get('ok')])"
for cfg in "--steps 40 --warmup 5" "--steps 40 --warmup 5 --ps_layout gpu0_ps_only --no_baseline" "--steps 40 --warmup 5 --ps_layout gpu0_ps_only --async_push 0 --no_baseline" "--model gpt2_medium --scheme ssgd_lr --steps 8 --warmup 3 --no_baseline" "--model bert_large --scheme tower_avg --steps 8 --warmup 3 --no_baseline"; do
  timeout 600 $RUN --master-port 29802 bench.py --gpus $N $cfg 2>/dev/null | tail -1 | python -c "
import sys, json
d = json.loads(sys.stdin.read()); print('$cfg', '->', '%.5g' % d['value'], '%.3f ms' % d['ms_per_step'], 'vs_baseline', d.get('vs_baseline'))"
done
