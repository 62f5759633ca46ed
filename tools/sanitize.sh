#!/bin/bash
# compute-sanitizer over every kernel family at small sizes.  ONE tool per gpurun call:
#     gpurun -- bash tools/sanitize.sh memcheck      (default)
#     gpurun -- bash tools/sanitize.sh racecheck
# The plain run must pass first; the log lands in gpurun_out/sanitize_<tool>.log and is copied
# to profiles/sanitize_<tool>.log when clean.  Hogwild's cross-GPU races are intentional
# (Hogwild/Hogwild.py:4-5) and are not what this checks: these are single-GPU kernels whose
# shared-memory / global-memory accesses must be race- and fault-free.
set -u
TOOL=${1:-memcheck}
O=gpurun_out
mkdir -p $O
python tools/sanitize_target.py > $O/sanitize_plain.log 2>&1 || { echo "plain run failed"; tail -30 $O/sanitize_plain.log; exit 1; }
tail -4 $O/sanitize_plain.log
cp $O/sanitize_plain.log $O/sanitize_selfcheck.log
EXTRA=""
[ "$TOOL" = memcheck ] && EXTRA="--leak-check no --padding 32"
[ "$TOOL" = racecheck ] && EXTRA="--racecheck-report all"
timeout 1500 compute-sanitizer --tool $TOOL $EXTRA --target-processes all --error-exitcode 7 \
    --log-file $O/sanitize_$TOOL.%p.log python tools/sanitize_target.py > $O/sanitize_${TOOL}_stdout.log 2>&1
rc=$?
echo "compute-sanitizer --tool $TOOL exit code $rc"
tail -5 $O/sanitize_${TOOL}_stdout.log
cat $O/sanitize_$TOOL.*.log > $O/sanitize_$TOOL.log 2>/dev/null; rm -f $O/sanitize_$TOOL.*.log
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|Error|Hazard|Invalid" $O/sanitize_$TOOL.log | sed 's/^=\+ *//; s/0x[0-9a-f]\+/ADDR/g' | sort | uniq -c | sort -rn | head -20
head -c 20000 $O/sanitize_$TOOL.log > $O/sanitize_${TOOL}_head.log
exit 0
