#!/bin/bash
# Round-2 evidence capture (run on the GPU box from the repo root):
#   gpurun --timeout 2400 -- 'bash profiles/capture_r2.sh'
# Plain runs first (ncu only after the same command exited 0 without it).
set -u
O=gpurun_out/r2
mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q --timeout 300 -p no:cacheprovider -rs > $O/pytest_gpu.log 2>&1; echo "pytest rc=$? $(tail -1 $O/pytest_gpu.log)"
timeout 600 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 5 --warmup 3 > $O/bench_reference.json 2>&1; echo "reference arm rc=$?"
timeout 600 python bench.py --suite gemv --fused --fmts q4_0,q4_1,q5_0,q5_1,q8_0,q2_k,q3_k,q4_k,q5_k,q6_k,f32,f16 > $O/sweep_decode_fused.jsonl 2>&1; echo "decode sweep rc=$?"
timeout 400 python bench.py --suite gemm --n 512 --fmts q4_0,q4_1,q5_0,q5_1,q8_0,q2_k,f32,f16,q3_k,q4_k,q5_k,q6_k > $O/sweep_gemm_n512.jsonl 2>&1; echo "gemm sweep rc=$?"
timeout 200 python bench.py --suite gemm --n 2048 --fmts q4_0,q4_1,q5_0,q5_1,q8_0,q2_k,f16 --shapes 5120x5120,13824x5120,5120x13824 > $O/sweep_gemm_13b_n2048.jsonl 2>&1
timeout 120 python bench.py --suite gemm --n 512 --fmts q5_0 --shapes 6656x6656,17920x6656,6656x17920 > $O/sweep_gemm_30b_q5_0_n512.jsonl 2>&1
timeout 90 python tools/trace_decode.py > $O/trace_decode.txt 2>&1
# launch list of the bench command (2 timed steps): per-launch durations + DRAM bytes, cold-cache and serialised
SHORT="--steps 2 --warmup 3 --no-cpu --no-e2e --no-extras --no-gemm --sustain-s 0"
timeout 300 python bench.py $SHORT > $O/bench_short.json 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none \
    -k regex:'decode|gemv|quantize|gemm|prep|wrow|xop|convert' -c 2500 --csv \
    --log-file $O/launches.csv python bench.py $SHORT > $O/ncu_launches.log 2>&1
echo "launch list rc=$?"
# full-section capture of the dominant kernel of the step, and of the two prefill kernels of the gemm block
timeout 600 ncu --set full --clock-control none --import-source on -k regex:decode16 -s 400 -c 6 -f -o $O/decode16_q41 \
    python bench.py $SHORT > $O/ncu_decode.log 2>&1
echo "decode capture rc=$?"
timeout 120 python tools/gemm_one.py q4_1 13824 5120 2048 > $O/gemm_one_q41.txt 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_so_kernel -s 2 -c 1 -f -o $O/gemm_so_pair_q41 \
    python tools/gemm_one.py q4_1 13824 5120 2048 > $O/ncu_gemm_so.log 2>&1
echo "gemm_so capture rc=$?"
ls -la $O
