#!/bin/bash
# Distributed-Setup: one ps and ONE worker (/root/reference/Distributed-Setup/run.sh, run_sup.sh).
#   ./run_setup.sh [--supervisor] [--fast ...]
exec "$(dirname "$0")/run.sh" distributed_setup.py --workers 1 "$@"
