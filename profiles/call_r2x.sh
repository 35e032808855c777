#!/bin/bash
# Round 2, call x: pair so-kernel with two unpack groups (768 threads)
set -u
O=gpurun_out/r2x
mkdir -p $O
timeout 90 python tools/gemm_one.py q4_1 4096 4096 512 > $O/first.txt 2>&1; rc=$?; echo "first rc=$rc"; tail -3 $O/first.txt
if [ $rc -ne 0 ]; then exit 0; fi
timeout 600 python -m pytest tests/test_gpu_gemm.py -m gpu -q -x --timeout 120 -p no:cacheprovider > $O/pytest_gemm.log 2>&1
echo "gemm tests rc=$? $(tail -1 $O/pytest_gemm.log)"; grep -E "^FAILED|^ERROR|Error|assert|err " $O/pytest_gemm.log | head -10
for f in q4_1 q4_0 q8_0 q5_1; do
for s in "13824 5120 2048" "11008 4096 512" "4096 4096 2048"; do
  LAMM_GEMM_PAIR=2 timeout 120 python tools/gemm_one.py $f $s 2>&1 | tail -1 | tee -a $O/gemm_pair.txt
  timeout 120 python tools/gemm_one.py $f $s 2>&1 | tail -1 | tee -a $O/gemm_auto.txt
done; done
