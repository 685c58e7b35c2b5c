#!/bin/bash
# ncu captures for profiles/ (run on the GPU box, one GPU): launch list of bench.py, and --set full captures of the
# search kernel (persistent form), the NSG link kernel and the filtered HNSW kernel. Each program has already exited 0
# without ncu in the same gpurun call.
set -x
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 1 > gpurun_out/b_plain.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_r1_bench_v6.csv \
    python bench.py --steps 2 --warmup 1 > gpurun_out/ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:beam_search_kernel -s 3 -c 1 -f -o gpurun_out/search_r1_v6 \
    python bench.py --steps 2 --warmup 1 > gpurun_out/ncu_search.log 2>&1
python scripts/nsg_phase_bench.py 200000 100 40 50 500 > gpurun_out/nsg_phase.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:link_kernel -c 1 -f -o gpurun_out/nsg_link_r1 \
    python scripts/nsg_phase_bench.py 200000 100 40 50 500 > gpurun_out/ncu_link.log 2>&1
python scripts/hnsw_filtered_bench.py 200000 128 16 100 100 > gpurun_out/filt.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:hnsw_filtered_kernel -s 3 -c 1 -f -o gpurun_out/hnsw_filtered_r1 \
    python scripts/hnsw_filtered_bench.py 200000 128 16 100 100 > gpurun_out/ncu_filt.log 2>&1
ls -la gpurun_out/
