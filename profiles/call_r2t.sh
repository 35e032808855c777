#!/bin/bash
# Round 2, call t: CTA-pair (cta_group::2) F16 prefill kernel -- hang guard first, then parity and rates
set -u
O=gpurun_out/r2t
mkdir -p $O
timeout 90 python tools/gemm_one.py f16 4096 4096 512 > $O/first.txt 2>&1; rc=$?; echo "first rc=$rc"; tail -3 $O/first.txt
if [ $rc -ne 0 ]; then exit 0; fi
timeout 300 python -m pytest tests/test_gpu_f16.py -m gpu -q -x --timeout 90 -p no:cacheprovider > $O/pytest_f16.log 2>&1
echo "f16 tests rc=$? $(tail -1 $O/pytest_f16.log)"; grep -E "^FAILED|^ERROR|Error|assert|err " $O/pytest_f16.log | head -10
for s in "4096 4096 512" "11008 4096 2048" "13824 5120 2048" "4096 11008 2048"; do
  timeout 120 python tools/gemm_one.py f16 $s 2>&1 | tail -1 | tee -a $O/gemm_f16_pair.txt
  LAMM_GEMM_PAIR=0 timeout 120 python tools/gemm_one.py f16 $s 2>&1 | tail -1 | tee -a $O/gemm_f16_single.txt
done
