#!/bin/bash
# ncu --set full of the tall pair kernel on two shapes + launch lists of a ResNet-50 and a GPT-2
# medium step (one GPU).  Summaries are produced on the box; big reports are dropped.
set -u
O=gpurun_out
mkdir -p $O
rm -f $O/*.ncu-rep
NCU="ncu --set full --clock-control none --import-source on --profile-from-start off"
summ() {
  python tools/ncu_summary.py $O/$1.ncu-rep "$2" > $O/$1_ncu_summary.md 2>> $O/summ.err
  ncu -i $O/$1.ncu-rep --page source --csv 2>> $O/summ.err | gzip -9 > $O/$1_source.csv.gz
  rm -f $O/$1.ncu-rep
}
python tools/gemm_prof.py 8192 8192 8192 > $O/p1.txt 2>&1 &&
python tools/gemm_prof.py 8192 4096 4096 > $O/p2.txt 2>&1 &&
python bench.py --model resnet50 --scheme adag --steps 1 --warmup 1 --update_window 1 --no_baseline --skip_e2e > $O/p3.txt 2>&1 &&
python bench.py --model gpt2_medium --scheme ssgd_lr --steps 1 --warmup 1 --no_baseline --skip_e2e > $O/p4.txt 2>&1 || { echo "plain run failed"; tail -20 $O/p?.txt; exit 1; }
cat $O/p1.txt $O/p2.txt; tail -c 300 $O/p3.txt; tail -c 300 $O/p4.txt
$NCU -o $O/gemm_tall_8k -f python tools/gemm_prof.py 8192 8192 8192 > $O/ncu1.log 2>&1
summ gemm_tall_8k "gemm_bf16_tall_kernel 8192^3 (tools/gemm_prof.py 8192 8192 8192)"
$NCU -o $O/gemm_tall_k4096 -f python tools/gemm_prof.py 8192 4096 4096 > $O/ncu2.log 2>&1
summ gemm_tall_k4096 "gemm_bf16_tall_kernel 8192x4096x4096"
ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file $O/resnet50_launches.csv \
    python bench.py --model resnet50 --scheme adag --steps 1 --warmup 1 --update_window 1 --no_baseline --skip_e2e > $O/ncu3.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file $O/gpt2_launches.csv \
    python bench.py --model gpt2_medium --scheme ssgd_lr --steps 1 --warmup 1 --no_baseline --skip_e2e > $O/ncu4.log 2>&1
ls -la $O | tail -20; du -sh $O
exit 0
