for ns in 0 500 2000 10000; do echo "== DTG_EPI_SLEEP_NS=$ns"; DTG_EPI_SLEEP_NS=$ns python tools/gemm_sweep.py --child 2>&1 | head -5; done
