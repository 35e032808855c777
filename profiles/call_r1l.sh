#!/bin/bash
set -u
O=gpurun_out/r1l
mkdir -p $O
timeout 400 python -m pytest tests -m gpu -q --timeout 200 -p no:cacheprovider -x > $O/pytest_gpu.log 2>&1; echo "pytest rc=$? $(tail -1 $O/pytest_gpu.log)"
grep -E "FAILED|ERROR" $O/pytest_gpu.log | head
timeout 300 python bench.py --suite gemm --n 512 --fmts q4_0,q4_1,q5_0,q5_1,q8_0,q2_k,f32,q3_k,q4_k,q5_k,q6_k > $O/sweep_gemm_n512.jsonl 2>&1; echo "gemm sweep rc=$?"
cut -c1-110 $O/sweep_gemm_n512.jsonl | head -30
timeout 120 python bench.py --suite gemm --n 2048 --fmts q4_1 --shapes 5120x5120,13824x5120,5120x13824 > $O/sweep_gemm_13b_n2048.jsonl 2>&1
timeout 120 python bench.py --suite gemm --n 512 --fmts q5_0 --shapes 6656x6656,17920x6656,6656x17920 > $O/sweep_gemm_30b_q5_0_n512.jsonl 2>&1
cat $O/sweep_gemm_13b_n2048.jsonl $O/sweep_gemm_30b_q5_0_n512.jsonl | cut -c1-110
timeout 300 python bench.py --suite gemv --fused --fmts q5_0,q5_1 > $O/sweep_decode_q5.jsonl 2>&1; echo "q5 sweep rc=$?"
cut -c1-130 $O/sweep_decode_q5.jsonl
