#!/bin/bash
# Round 2, call v: ncu --set full of the CTA-pair scaled-operand GEMM (13B Q4_1 N=2048), after the same command ran plain
set -u
O=gpurun_out/r2v
mkdir -p $O
timeout 120 python tools/gemm_one.py q4_1 13824 5120 2048 > $O/plain.txt 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_so_kernel -s 2 -c 1 -o $O/so_pair -f python tools/gemm_one.py q4_1 13824 5120 2048 > $O/ncu.log 2>&1
echo "rc=$?"; tail -2 $O/plain.txt; tail -3 $O/ncu.log; ls -la $O
