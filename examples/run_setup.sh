#!/bin/bash
# One ps and ONE worker (reference: Distributed-Setup/run.sh and run_sup.sh).
# Usage: ./run_setup.sh [--supervisor] [--fast ...]
here=$(dirname "$0")
python "$here/distributed_setup.py" --job_name "ps" --task_index 0 "$@" &
python "$here/distributed_setup.py" --job_name "worker" --task_index 0 "$@" &
wait
