TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29616"
for nv in 1 0; do
echo -n "DTG_NVLS=$nv "
DTG_NVLS=$nv timeout 200 $TR bench.py --gpus 4 --model gpt2_medium --scheme ssgd_lr --steps 8 --warmup 3 --skip_e2e 2>&1 | grep '^{\|Error\|error' | grep -o "\"ms_per_step\": [0-9.]*\|.*rror.*\|fused reduce[^\"]*" | tr "\n" " "; echo
done
