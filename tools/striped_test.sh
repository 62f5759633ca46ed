TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29613"
timeout 200 $TR bench.py --gpus 8 --steps 50 --warmup 5 --ps_layout striped 2>&1 | grep '^{\|Error\|error' | cut -c1-300
timeout 200 $TR bench.py --gpus 8 --steps 50 --warmup 5 2>&1 | grep '^{\|Error\|error' | cut -c1-300
timeout 200 python -m pytest tests/test_gpu_engine.py -m gpu -x -q -k "striped" 2>&1 | tail -3
TR4="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29614"
timeout 200 $TR4 bench.py --gpus 4 --steps 50 --warmup 5 --ps_layout striped 2>&1 | grep '^{\|Error\|error' | cut -c1-300
