#!/bin/bash
set -u
O=gpurun_out/r1d
mkdir -p $O
timeout 600 python -m pytest tests -m gpu -q --timeout 200 -p no:cacheprovider > $O/pytest_gpu.log 2>&1; echo "pytest rc=$? $(tail -1 $O/pytest_gpu.log)"
grep -E "FAILED|ERROR" $O/pytest_gpu.log | head -30
timeout 90 python tools/trace_decode.py > $O/trace_decode.txt 2>&1; echo "decode trace rc=$?"
timeout 90 python tools/trace_decode.py --fmt q8_0 >> $O/trace_decode.txt 2>&1
timeout 90 python tools/trace_decode.py --fmt q4_k >> $O/trace_decode.txt 2>&1
cat $O/trace_decode.txt
