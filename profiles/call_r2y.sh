#!/bin/bash
# Round 2, call y: ring-depth variants of the pair so-kernel (A stages / B stages), Q4_0 and Q4_1 at 13B N=2048
set -u
O=gpurun_out/r2y
mkdir -p $O
for v in default 2_3 3_3 4_2; do
  for f in q4_0 q4_1; do
    if [ $v = default ]; then L=""; else L="$PWD/build_variants/liblammcu_$v.so"; fi
    echo -n "$v $f: " | tee -a $O/variants.txt
    LAMM_LIB=$L LAMM_GEMM_PAIR=2 timeout 120 python tools/gemm_one.py $f 13824 5120 2048 2>&1 | tail -1 | tee -a $O/variants.txt
  done
done
