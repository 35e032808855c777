#!/bin/bash
# Round 2, call ba: tensor peaks again, now with the CTA-pair (cta_group::2) M256 x N256 fp16 instruction
set -u
O=gpurun_out/r2ba
mkdir -p $O
nvcc -gencode arch=compute_100a,code=sm_100a -O2 -Wno-deprecated-gpu-targets -o $O/i8peak tools/i8peak.cu 2>&1 | tail -2
timeout 120 $O/i8peak > $O/i8peak.txt 2>&1; echo rc=$?; cat $O/i8peak.txt; rm -f $O/i8peak
