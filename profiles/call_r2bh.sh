#!/bin/bash
# Round 2, call bh: raw chunk pitch padded for Q4_1 / Q5_1 (no ld.shared bank conflicts) -- GEMM suites, rates
set -u
O=gpurun_out/r2bh
mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_gemm.py tests/test_gpu_fullsize.py -m gpu -q --timeout 300 -p no:cacheprovider > $O/pytest_gemm.log 2>&1
echo "gemm tests rc=$? $(tail -1 $O/pytest_gemm.log)"; grep -E "^FAILED|^ERROR" $O/pytest_gemm.log | head -10
for f in q4_1 q5_1; do
for a in 0 1; do
  echo "LAMM_GEMM_ATM=$a" | tee -a $O/gemm_rates.txt
  LAMM_GEMM_ATM=$a timeout 120 python tools/gemm_one.py $f 13824 5120 2048 2>&1 | tail -1 | tee -a $O/gemm_rates.txt
  LAMM_GEMM_ATM=$a timeout 120 python tools/gemm_one.py $f 11008 4096 512 2>&1 | tail -1 | tee -a $O/gemm_rates.txt
done; done
LAMM_GEMM_PAIR=0 timeout 120 python tools/gemm_one.py q4_1 4096 4096 512 2>&1 | tail -1 | tee -a $O/gemm_rates.txt
LAMM_GEMM_PAIR=0 timeout 120 python tools/gemm_one.py q5_1 4096 4096 512 2>&1 | tail -1 | tee -a $O/gemm_rates.txt
timeout 120 python tools/gemm_sustained.py q4_1 13824 5120 2048 2 2>&1 | tail -1 | tee -a $O/gemm_rates.txt
