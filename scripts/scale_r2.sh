# one N of the scaling run, the way the driver launches it
N=${1:-8}
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 100 --warmup 4 > gpurun_out/bench_r2_n$N.json 2> gpurun_out/bench_r2_n$N.err
echo "rc=$?"; tail -4 gpurun_out/bench_r2_n$N.err; head -c 1500 gpurun_out/bench_r2_n$N.json; echo
