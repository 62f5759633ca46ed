#!/bin/bash
# One ps and two workers on localhost, each its own process, like the reference's run.sh
# files.  Usage: ./run.sh hogwild.py [--fast ...]
script=${1:-hogwild.py}; shift
python "$script" --job_name "ps" --task_index 0 "$@" &
python "$script" --job_name "worker" --task_index 0 "$@" &
python "$script" --job_name "worker" --task_index 1 "$@" &
wait
