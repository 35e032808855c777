#!/bin/bash
# Round-1 evidence capture (run on the GPU box from the repo root):
#   gpurun --timeout 1500 -- 'bash profiles/capture_r1.sh'
# Plain runs first (ncu only after the same command exited 0 without it).
set -u
O=gpurun_out/r1
mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"
python bench.py --steps 200 --warmup 5 > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"
python bench.py --suite gemv --fused > $O/sweep_decode_fused.jsonl 2>&1; echo "sweep rc=$?"
python bench.py --suite gemm --n 512 > $O/sweep_gemm_n512.jsonl 2>&1; echo "gemm sweep rc=$?"
python bench.py --suite gemm --n 2048 --fmts q4_1 --shapes 5120x5120,13824x5120,5120x13824 > $O/sweep_gemm_13b_n2048.jsonl 2>&1
# launch list of the bench command (2 timed steps): per-launch durations, cold-cache and serialised
python bench.py --steps 2 --warmup 3 --no-cpu > $O/bench_short.json 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'decode|gemv|quantize|gemm' -c 2500 --csv \
    --log-file $O/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu > $O/ncu_launches.log 2>&1
echo "launch list rc=$?"
# full-section captures of the dominant kernels
ncu --set full --clock-control none --import-source on -k regex:decode16 -s 450 -c 8 -f -o $O/decode16_q41 \
    python bench.py --suite gemv --fused --fmts q4_1 --shapes 4096x4096,11008x4096,4096x11008,32000x4096 --ring-mb 64 > $O/ncu_decode.log 2>&1
echo "decode capture rc=$?"
ncu --set full --clock-control none --import-source on -k regex:gemm_tc -s 4 -c 1 -f -o $O/gemm_tc_q41 \
    python bench.py --suite gemm --n 512 --fmts q4_1 --shapes 4096x4096 --ring-mb 64 > $O/ncu_gemm.log 2>&1
echo "gemm capture rc=$?"
ls -la $O
