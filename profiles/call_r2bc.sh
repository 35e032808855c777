#!/bin/bash
# Round 2, call bc: pair GEMM with the unpacked weight tiles in TENSOR MEMORY (TS-form MMAs) -- hang guard, parity, rates
set -u
O=gpurun_out/r2bc
mkdir -p $O
timeout 90 python tools/gemm_one.py q4_1 4096 4096 512 > $O/first.txt 2>&1; rc=$?; echo "first rc=$rc"; tail -3 $O/first.txt
if [ $rc -ne 0 ]; then exit 0; fi
timeout 120 python tools/sanitize_small.py 2,3,6,7,8 > $O/small.txt 2>&1; echo "small rc=$? $(tail -1 $O/small.txt)"
for f in q4_1 q4_0 q5_1 q8_0; do
for s in "13824 5120 2048" "11008 4096 512"; do
  timeout 120 python tools/gemm_one.py $f $s 2>&1 | tail -1 | tee -a $O/gemm_atm.txt
  LAMM_GEMM_ATM=0 timeout 120 python tools/gemm_one.py $f $s 2>&1 | tail -1 | tee -a $O/gemm_smem.txt
done; done
timeout 900 python -m pytest tests/test_gpu_gemm.py tests/test_gpu_fullsize.py -m gpu -q -x --timeout 200 -p no:cacheprovider > $O/pytest_gemm.log 2>&1
echo "gemm tests rc=$? $(tail -1 $O/pytest_gemm.log)"; grep -E "^FAILED|^ERROR|Error|assert|err " $O/pytest_gemm.log | head -10
