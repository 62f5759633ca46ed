L=distributed-tensorflow-guide_b200/ops/libdtg_b200.so
for rep in 1 2; do for v in before after; do
cp tools/_ab/libdtg_$v.so $L
echo -n "$v "; timeout 100 python tools/gemm_prof.py 8192 8192 8192 | tr "\n" " "; timeout 100 python tools/gemm_prof.py 8192 4096 4096 | tr "\n" " "
timeout 200 python bench.py --steps 100 --warmup 10 --skip_e2e 2>&1 | tail -1 | grep -o "\"ms_per_step\": [0-9.]*"
done; done
cp tools/_ab/libdtg_before.so $L
timeout 300 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q 2>&1 | tail -2
