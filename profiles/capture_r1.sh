#!/bin/bash
# Round-1 evidence capture (run on the GPU box from the repo root):
#   gpurun --timeout 1500 -- 'bash profiles/capture_r1.sh'
# Plain runs first (ncu only after the same command exited 0 without it).
set -u
O=gpurun_out/r1
mkdir -p $O
timeout 400 python -m pytest tests -m gpu -q --timeout 200 -p no:cacheprovider > $O/pytest_gpu.log 2>&1; echo "pytest rc=$? $(tail -1 $O/pytest_gpu.log)"
timeout 300 python bench.py --steps 200 --warmup 5 > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"
timeout 200 python bench.py --steps 200 --warmup 5 --no-batch --no-cpu > $O/bench_nobatch.json 2>&1; echo "bench (one launch per matmul) rc=$?"
timeout 300 python bench.py --impl reference --steps 5 --warmup 1 > $O/bench_reference.json 2>&1; echo "reference arm rc=$?"
timeout 400 python bench.py --suite gemv --fused --fmts q4_0,q4_1,q5_0,q5_1,q8_0,q2_k,q3_k,q4_k,q5_k,q6_k,f32 > $O/sweep_decode_fused.jsonl 2>&1; echo "sweep rc=$?"
timeout 300 python bench.py --suite gemm --n 512 --fmts q4_0,q4_1,q5_0,q5_1,q8_0,q2_k,f32,q3_k,q4_k,q5_k,q6_k > $O/sweep_gemm_n512.jsonl 2>&1; echo "gemm sweep rc=$?"
timeout 120 python bench.py --suite gemm --n 2048 --fmts q4_1 --shapes 5120x5120,13824x5120,5120x13824 > $O/sweep_gemm_13b_n2048.jsonl 2>&1
timeout 120 python bench.py --suite gemm --n 512 --fmts q5_0 --shapes 6656x6656,17920x6656,6656x17920 > $O/sweep_gemm_30b_q5_0_n512.jsonl 2>&1
LAMM_SERVER_TRACE=1 timeout 90 python tools/trace_server.py > $O/trace_server.txt 2>&1
LAMM_SERVER_TRACE=1 timeout 90 python tools/trace_server.py --m 4096 --k 11008 >> $O/trace_server.txt 2>&1
timeout 90 python tools/trace_decode.py > $O/trace_decode.txt 2>&1
nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o $O/hostlat tools/hostlat.cu 2>/dev/null && timeout 60 $O/hostlat > $O/hostlat.txt 2>&1; rm -f $O/hostlat
# launch list of the bench command (2 timed steps): per-launch durations + DRAM bytes, cold-cache and serialised
timeout 200 python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e > $O/bench_short.json 2>&1 && \
timeout 400 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none \
    -k regex:'decode|gemv|quantize|gemm' -c 1500 --csv \
    --log-file $O/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e > $O/ncu_launches.log 2>&1
echo "launch list rc=$?"
# full-section capture of the dominant kernel
timeout 400 ncu --set full --clock-control none --import-source on -k regex:decode16 -s 400 -c 6 -f -o $O/decode16_q41 \
    python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e > $O/ncu_decode.log 2>&1
echo "decode capture rc=$?"
ls -la $O
