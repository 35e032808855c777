#!/bin/bash
set -u
O=gpurun_out/r1m
mkdir -p $O
timeout 300 python -m pytest tests -m gpu -q --timeout 200 -p no:cacheprovider -x > $O/pytest_gpu.log 2>&1; echo "pytest rc=$? $(tail -1 $O/pytest_gpu.log)"
grep -E "FAILED|ERROR" $O/pytest_gpu.log | head
timeout 100 python bench.py --suite gemv --fused --fmts q8_0 --shapes 4096x4096,11008x4096,4096x11008,5120x5120,6656x6656,17920x6656 > $O/sweep_q8.jsonl 2>&1
cut -c1-130 $O/sweep_q8.jsonl
