#!/bin/bash
# Round 2, call bo: persistent F16 pair GEMM (double-buffered TMEM accumulators) -- hang guard, F16 tests, rates
set -u
O=gpurun_out/r2bo
mkdir -p $O
timeout 90 python tools/gemm_one.py f16 4096 4096 512 > $O/first.txt 2>&1; rc=$?; echo "first rc=$rc"; tail -1 $O/first.txt
if [ $rc -ne 0 ]; then tail -5 $O/first.txt; exit 0; fi
timeout 600 python -m pytest tests/test_gpu_f16.py -m gpu -q -x --timeout 200 -p no:cacheprovider > $O/pytest_f16.log 2>&1
echo "f16 tests rc=$? $(tail -1 $O/pytest_f16.log)"; grep -E "^FAILED|^ERROR|assert " $O/pytest_f16.log | head -5
for s in "13824 5120 2048" "5120 13824 2048" "11008 4096 512" "13824 5120 512" "4096 4096 2048"; do
  timeout 120 python tools/gemm_one.py f16 $s 2>&1 | tail -1 | tee -a $O/gemm_f16.txt
done
timeout 120 python tools/gemm_sustained.py f16 13824 5120 2048 2 2>&1 | tail -1 | tee -a $O/gemm_f16.txt
