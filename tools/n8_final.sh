#!/bin/bash
# Last 8-GPU pass of round 2: NVLink micro-benchmarks (hot spot, NVLS vs peer), then the configs
# whose kernels changed after the matrix run.
set -u
N=${1:-8}
O=gpurun_out/final_n$N
mkdir -p $O
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 400 $RUN --master-port 29851 tools/nvlink_bench.py 2>&1 | grep -E "^(NVLINK|HOTSPOT|SYNCBW)" | tee $O/nvlink.txt | cut -c1-700
p=29860
for cfg in "hogwild_striped|--steps 40 --warmup 5" "resnet_adag|--model resnet50 --scheme adag --steps 3 --warmup 2" "gpt2_ssgd_lr|--model gpt2_medium --scheme ssgd_lr --steps 8 --warmup 3" "bert_tower_avg|--model bert_large --scheme tower_avg --steps 8 --warmup 3"; do
  name=${cfg%%|*}; args=${cfg#*|}; p=$((p+1))
  timeout 900 $RUN --master-port $p bench.py --gpus $N $args 2> $O/$name.err | tail -1 > $O/$name.json
  python -c "
import json;d=json.load(open('$O/$name.json'));b=d.get('baseline') or {}
print('%-16s N=%d native %.5g (%.3f ms) e2e %.5g | torch %.5g (%.3f ms) | ratio %.3f' % ('$name', d['n_gpus'], d['value'], d['ms_per_step'], d['e2e']['value'], b.get('value',0), b.get('ms_per_step',0), d.get('vs_baseline') or 0))"
done
exit 0
