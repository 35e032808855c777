#!/bin/bash
# TS-form (A in TMEM) known answer + rate
mkdir -p gpurun_out/r2bb
nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o /tmp/tmema tools/tmema.cu > gpurun_out/r2bb/build.log 2>&1
timeout 60 /tmp/tmema 4000 > gpurun_out/r2bb/tmema.log 2>&1; echo "rc=$?" >> gpurun_out/r2bb/tmema.log
cat gpurun_out/r2bb/tmema.log
