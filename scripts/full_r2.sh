# round-2 full-size evidence run on one B200: GPU tests, both bench arms, launch list, ncu of the benchmarked search launch
python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests_r2_full.log 2>&1; tail -4 gpurun_out/gpu_tests_r2_full.log
timeout 900 python bench.py --impl reference --steps 5 --warmup 3 > gpurun_out/bench_ref_r2.json 2> gpurun_out/bench_ref_r2.err; echo "ref rc=$?"; cat gpurun_out/bench_ref_r2.json | head -c 600; echo
timeout 1200 python bench.py > gpurun_out/bench_r2.json 2> gpurun_out/bench_r2.err; echo "ours rc=$?"; tail -5 gpurun_out/bench_r2.err
timeout 300 python scripts/exp_search_r2.py --only ncu --L 32 > gpurun_out/ncu_plain.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:beam_search2_kernel -s 1 -c 1 -o gpurun_out/search_r2_final python scripts/exp_search_r2.py --only ncu --L 32 > gpurun_out/ncu_run.log 2>&1; tail -2 gpurun_out/ncu_run.log
