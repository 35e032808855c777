#!/bin/bash
# Round 2, call bd: burst vs SUSTAINED tensor rates (power cap): pure MMA stream, and the pair GEMM with A in shared memory / in TMEM
set -u
O=gpurun_out/r2bd
mkdir -p $O
nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o /tmp/i8peak tools/i8peak.cu > $O/build.log 2>&1
timeout 120 /tmp/i8peak 4000 3 2>&1 | tee $O/i8peak_sustained.txt
for rep in 1 2; do
for a in 1 0; do
  echo "LAMM_GEMM_ATM=$a" | tee -a $O/gemm_sustained.txt
  LAMM_GEMM_ATM=$a timeout 120 python tools/gemm_sustained.py q4_1 13824 5120 2048 3 2>&1 | tail -1 | tee -a $O/gemm_sustained.txt
done; done
LAMM_GEMM_ATM=1 timeout 120 python tools/gemm_sustained.py q2_k 13824 5120 2048 3 2>&1 | tail -1 | tee -a $O/gemm_sustained.txt
LAMM_GEMM_ATM=1 timeout 120 python tools/gemm_sustained.py f16 13824 5120 2048 3 2>&1 | tail -1 | tee -a $O/gemm_sustained.txt
