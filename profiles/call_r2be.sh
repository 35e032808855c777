#!/bin/bash
# Round 2, call be: every form of the scaled-operand GEMM on ragged shapes (new test), Q5_0 / Q5_1 with A in TMEM vs shared memory
set -u
O=gpurun_out/r2be
mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_gemm.py -m gpu -q -x --timeout 300 -p no:cacheprovider -k "every_form" > $O/pytest_forms.log 2>&1
echo "forms rc=$? $(tail -1 $O/pytest_forms.log)"; grep -E "^FAILED|^ERROR|Error|assert|err " $O/pytest_forms.log | head -10
for f in q5_0 q5_1; do
for a in 1 0; do
  echo "LAMM_GEMM_ATM=$a" | tee -a $O/gemm_sustained.txt
  LAMM_GEMM_ATM=$a timeout 120 python tools/gemm_sustained.py $f 13824 5120 2048 2 2>&1 | tail -1 | tee -a $O/gemm_sustained.txt
  LAMM_GEMM_ATM=$a timeout 120 python tools/gemm_one.py $f 11008 4096 512 2>&1 | tail -1 | tee -a $O/gemm_sustained.txt
done; done
timeout 900 python -m pytest tests/test_gpu_gemm.py tests/test_gpu_fullsize.py tests/test_gpu_f16.py -m gpu -q -x --timeout 300 -p no:cacheprovider > $O/pytest_gemm.log 2>&1
echo "gemm tests rc=$? $(tail -1 $O/pytest_gemm.log)"; grep -E "^FAILED|^ERROR|Error|assert|err " $O/pytest_gemm.log | head -10
