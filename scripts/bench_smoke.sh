# small-size shake-out of bench.py (both arms, N=1 and N=2 when two GPUs are visible)
export B200ANN_BENCH_N=${B200ANN_BENCH_N:-200000} B200ANN_BENCH_C5_N=${B200ANN_BENCH_C5_N:-100000} B200ANN_BENCH_CPU_BUILD_ROWS=20000
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bs_ours1.json 2> gpurun_out/bs_ours1.err; echo "ours rc=$?"; tail -3 gpurun_out/bs_ours1.err; head -c 3000 gpurun_out/bs_ours1.json; echo
timeout 600 python bench.py --impl reference --steps 3 --warmup 3 > gpurun_out/bs_ref1.json 2> gpurun_out/bs_ref1.err; echo "ref rc=$?"; tail -2 gpurun_out/bs_ref1.err; head -c 1500 gpurun_out/bs_ref1.json; echo
if [ "$(nvidia-smi -L | wc -l)" -ge 2 ]; then
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bs_ours2.json 2> gpurun_out/bs_ours2.err; echo "ours2 rc=$?"; tail -3 gpurun_out/bs_ours2.err; head -c 3000 gpurun_out/bs_ours2.json; echo
fi
