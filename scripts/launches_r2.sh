# round-2 launch list of the bench command (one GPU): plain run first, then the same command under ncu
timeout 900 python bench.py --no-extras --steps 2 --warmup 1 > gpurun_out/bench_short_r2.json 2> gpurun_out/bench_short_r2.err; echo "plain rc=$?"
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_r2_bench.csv python bench.py --no-extras --steps 2 --warmup 1 > gpurun_out/ncu_bench_r2.log 2>&1; echo "ncu rc=$?"
tail -3 gpurun_out/ncu_bench_r2.log; wc -l gpurun_out/launches_r2_bench.csv
