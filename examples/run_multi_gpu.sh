#!/bin/bash
# ps hidden from the GPUs, worker k pinned to GPU k
# (reference: Multiple-GPUs-Single-Machine/dist_mult_gpu_sing_mach.sh).
CUDA_VISIBLE_DEVICES=-1 python hogwild.py --gpu --job_name "ps" --task_index 0 "$@" &
python hogwild.py --gpu --job_name "worker" --task_index 0 "$@" &
python hogwild.py --gpu --job_name "worker" --task_index 1 "$@" &
wait
