# round-2 closing evidence run on one B200: GPU tests, both bench arms, launch list, the side benches quoted in README / BASELINE.md
python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests_r2_full.log 2>&1; tail -3 gpurun_out/gpu_tests_r2_full.log
timeout 900 python bench.py --impl reference --steps 5 --warmup 3 > gpurun_out/bench_ref_r2.json 2> gpurun_out/bench_ref_r2.err; echo "ref rc=$?"
timeout 1200 python bench.py > gpurun_out/bench_r2.json 2> gpurun_out/bench_r2.err; echo "ours rc=$?"
bash scripts/launches_r2.sh
timeout 600 python scripts/run_configs.py c1 c3 c4 > gpurun_out/configs_r2.jsonl 2> gpurun_out/configs_r2.err; echo "configs rc=$?"; cut -c1-400 gpurun_out/configs_r2.jsonl
timeout 300 python scripts/hnsw_live_bench.py > gpurun_out/hnsw_live.log 2>&1; echo "live rc=$?"
timeout 300 python scripts/hnsw_u8_bench.py > gpurun_out/hnsw_u8.log 2>&1; echo "u8 rc=$?"; tail -3 gpurun_out/hnsw_u8.log | cut -c1-300
