#!/bin/bash
# Final evidence pass of round 2 (one GPU): full GPU test suite, kernel zoo timings + ncu --set
# full summaries, launch lists of a ResNet-50 and a GPT-2 medium step, bench line.
set -u
O=gpurun_out
mkdir -p $O
rm -f $O/*.ncu-rep
python -m pytest tests -m gpu -x -q 2>&1 | tail -6
NCU="ncu --set full --clock-control none --import-source on --profile-from-start off"
python tools/kernel_zoo.py > $O/kernel_zoo_stdout.txt 2>&1 &&
python bench.py --model resnet50 --scheme adag --steps 1 --warmup 1 --update_window 1 --no_baseline --skip_e2e > $O/p3.txt 2>&1 &&
python bench.py --model gpt2_medium --scheme ssgd_lr --steps 1 --warmup 1 --no_baseline --skip_e2e > $O/p4.txt 2>&1 || { echo "plain run failed"; tail -20 $O/kernel_zoo_stdout.txt $O/p3.txt $O/p4.txt; exit 1; }
tail -52 $O/kernel_zoo_stdout.txt
$NCU -o $O/kernel_zoo -f python tools/kernel_zoo.py > $O/ncu5.log 2>&1
python tools/ncu_summary.py $O/kernel_zoo.ncu-rep "every non-GEMM kernel family + implicit-GEMM convolutions, one launch each (tools/kernel_zoo.py)" > $O/kernel_zoo_ncu_summary.md 2>> $O/summ.err
rm -f $O/kernel_zoo.ncu-rep
ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file $O/resnet50_launches.csv \
    python bench.py --model resnet50 --scheme adag --steps 1 --warmup 1 --update_window 1 --no_baseline --skip_e2e > $O/ncu3.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file $O/gpt2_launches.csv \
    python bench.py --model gpt2_medium --scheme ssgd_lr --steps 1 --warmup 1 --no_baseline --skip_e2e > $O/ncu4.log 2>&1
python bench.py --steps 60 --warmup 10 2>/dev/null | tail -1 > $O/bench_n1.json; python -c "
import json;d=json.load(open('$O/bench_n1.json'));print('hogwild N=1',d['value'],d['ms_per_step'],d['vs_baseline'],d['e2e']['value'])"
du -sh $O
exit 0
