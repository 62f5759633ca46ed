set -x
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29611"
timeout 600 python -m pytest tests/test_gpu_engine.py -m gpu -x -q 2>&1 | tail -3
for impl in native torch; do
timeout 300 $TR bench.py --gpus 8 --steps 50 --warmup 5 --impl $impl 2>&1 | grep '^{' | cut -c1-220
timeout 300 $TR bench.py --gpus 8 --model gpt2_medium --scheme ssgd_lr --steps 10 --warmup 3 --impl $impl 2>&1 | grep '^{' | cut -c1-220
timeout 300 $TR bench.py --gpus 8 --model bert_large --scheme tower_avg --steps 10 --warmup 3 --impl $impl 2>&1 | grep '^{' | cut -c1-220
timeout 400 $TR bench.py --gpus 8 --model resnet50 --scheme adag --steps 4 --warmup 3 --impl $impl 2>&1 | grep '^{' | cut -c1-220
done
