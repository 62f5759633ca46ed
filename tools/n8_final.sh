TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29617"
timeout 200 $TR bench.py --gpus 8 --steps 50 --warmup 5 2>&1 | grep '^{\|Error\|error' | cut -c1-1800
timeout 200 $TR bench.py --gpus 8 --model gpt2_medium --scheme ssgd_lr --steps 10 --warmup 3 2>&1 | grep '^{\|Error\|error' | cut -c1-1800
