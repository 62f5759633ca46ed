#!/bin/bash
# Every role of a localhost cluster from one command: one ps, then N workers (N = the value of a
# `--workers N` flag if given, else 2), each its own OS process -- what the reference's run.sh
# files do by hand for N = 2 (/root/reference/Hogwild/run.sh).
#   ./run.sh hogwild.py [--fast] [--workers 3] [script flags...]
set -u
here=$(cd "$(dirname "$0")" && pwd)
script=${1:-hogwild.py}
[ $# -gt 0 ] && shift
[ -f "$script" ] || script="$here/$script"      # a path, or the name of an example next to this file
n_workers=2
args=("$@")
for ((i = 0; i < ${#args[@]}; i++)); do
  [ "${args[$i]}" = "--workers" ] && n_workers=${args[$((i + 1))]}
done
launch() {  # job, index
  python "$script" --job_name "$1" --task_index "$2" "${args[@]}" &
  pids+=($!)
}
pids=()
launch ps 0
for ((k = 0; k < n_workers; k++)); do launch worker "$k"; done
rc=0
for p in "${pids[@]}"; do wait "$p" || rc=$?; done
exit $rc
