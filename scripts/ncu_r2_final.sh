#!/bin/bash
# closing ncu captures of round 2 (one GPU): the kNN sweep's main pass after the folded norm, and one full-size batch of the
# HNSW insertion kernel after the one-pass pool merge. Each program exits 0 without ncu first.
mkdir -p gpurun_out
STEPS=16 timeout 200 python scripts/knn_sample_ab.py > gpurun_out/ncu_knn_plain.log 2>&1 || { echo "knn plain failed"; exit 1; }
STEPS=16 timeout 900 ncu --set full --clock-control none --import-source on -k regex:sweep_kernel -s 1 -c 1 -f -o gpurun_out/knn_sweep_r2_final \
    python scripts/knn_sample_ab.py > gpurun_out/ncu_knn.log 2>&1; echo "ncu knn rc=$?"
SIZES=1000000 RATIOS=32 timeout 200 python scripts/hnsw_ratio_ab.py > gpurun_out/ncu_hnsw_plain.log 2>&1 || { echo "hnsw plain failed"; exit 1; }
SIZES=1000000 RATIOS=32 timeout 900 ncu --set full --clock-control none --import-source on -k regex:hnsw_insert_kernel -s 380 -c 1 -f -o gpurun_out/hnsw_insert_r2_final \
    python scripts/hnsw_ratio_ab.py > gpurun_out/ncu_hnsw.log 2>&1; echo "ncu hnsw rc=$?"
ls -la gpurun_out/*.ncu-rep
