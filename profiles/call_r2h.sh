#!/bin/bash
# Round 2, call h: the no-allocation entry point's test alone in a fresh process, then after the whole GPU suite (-x).
set -u
O=gpurun_out/r2h
mkdir -p $O
timeout 200 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x --timeout 120 -p no:cacheprovider > $O/pytest_alone.log 2>&1
echo "alone rc=$? $(tail -1 $O/pytest_alone.log)"; grep -E "^FAILED|^ERROR|Error" $O/pytest_alone.log | head
timeout 900 python -m pytest tests -m gpu -q -x --timeout 300 -p no:cacheprovider > $O/pytest_gpu.log 2>&1
echo "suite rc=$? $(tail -1 $O/pytest_gpu.log)"; grep -E "^FAILED|^ERROR" $O/pytest_gpu.log | head
