timeout 300 python -m pytest tests/test_gpu_search_v2.py -x -q > gpurun_out/gpu_tests_v2.log 2>&1; tail -3 gpurun_out/gpu_tests_v2.log
for tail in -1 1184 2368 3552; do echo "TAIL $tail"; B200ANN_SEARCH_TAIL=$tail timeout 300 python scripts/exp_search_r2.py --L 32 --only quick 2>&1 | grep '"tag"' ; done
