#!/bin/bash
# Round 2, call bj (2 GPUs): the multi-GPU test + smoke + a 2-GPU bench line at the final kernels (raw pad, dependent arrive, server rings)
set -u
O=gpurun_out/r2bj
mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -q --timeout 800 -p no:cacheprovider -rs > $O/pytest_multi.log 2>&1; echo "multi rc=$? $(tail -1 $O/pytest_multi.log)"
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.txt 2>&1; echo "smoke rc=$?"; tail -4 $O/smoke.txt
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 3 > $O/bench_n2.json 2> $O/bench_n2.err; echo "bench n2 rc=$?"
tail -c 1500 $O/bench_n2.json
