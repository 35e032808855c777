#!/bin/bash
# 2-GPU parity + bench (gpurun --gpus 2): decode LL exchange, sharded sibling batches, row-sharded prefill
set -u
O=gpurun_out/mg2
mkdir -p $O
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 \
  tests/multi_gpu_check.py > $O/check_n2.log 2>&1; echo "check n=2 rc=$?"
grep -E "ok|FAIL" $O/check_n2.log | tail -30
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29535 \
  bench.py --gpus 2 --steps 100 --warmup 5 --no-cpu > $O/bench_n2.json 2> $O/bench_n2.err; echo "bench n=2 rc=$? $(tail -1 $O/bench_n2.json | cut -c1-200)"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29536 \
  bench.py --gpus 2 --steps 100 --warmup 5 --no-cpu --model 30b --fmt q5_0 > $O/bench_30b_n2.json 2> $O/bench_30b_n2.err; echo "bench 30b n=2 rc=$? $(tail -1 $O/bench_30b_n2.json | cut -c1-200)"
