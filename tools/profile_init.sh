#!/bin/bash
# Timing + one ncu --set full capture of the device-side initialiser kernel (one GPU).
O=gpurun_out/init; mkdir -p $O
python tools/kernel_zoo.py --only init > $O/zoo.txt 2>&1 || { echo "plain run failed"; tail -20 $O/zoo.txt; exit 1; }
cat $O/zoo.txt | tail -5
ncu --set full --clock-control none --import-source on --profile-from-start off -o $O/init -f \
    python tools/kernel_zoo.py --only init > $O/ncu.log 2>&1
python tools/ncu_summary.py $O/init.ncu-rep "device-side initialisers, 51.5 M elements (tools/kernel_zoo.py --only init)" > $O/init_ncu_summary.md 2>> $O/summ.err
rm -f $O/init.ncu-rep
head -40 $O/init_ncu_summary.md
