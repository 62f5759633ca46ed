#!/bin/bash
# Multiple GPUs, single machine: the ps process sees no GPU, worker k sees only GPU k
# (/root/reference/Multiple-GPUs-Single-Machine/dist_mult_gpu_sing_mach.sh does this with
# CUDA_VISIBLE_DEVICES for two workers).   ./run_multi_gpu.sh [--workers N] [flags...]
set -u
here=$(cd "$(dirname "$0")" && pwd)
n_workers=2
args=("$@")
for ((i = 0; i < ${#args[@]}; i++)); do
  [ "${args[$i]}" = "--workers" ] && n_workers=${args[$((i + 1))]}
done
pids=()
role() {  # visible devices, job, index
  CUDA_VISIBLE_DEVICES=$1 python "$here/hogwild.py" --gpu --job_name "$2" --task_index "$3" "${args[@]}" &
  pids+=($!)
}
role -1 ps 0
for ((k = 0; k < n_workers; k++)); do role "$k" worker "$k"; done
rc=0
for p in "${pids[@]}"; do wait "$p" || rc=$?; done
exit $rc
