#!/bin/bash
# Round-1 multi-GPU scaling capture.  Charged N x box time, so each N runs on a box of its own size:
#   gpurun --gpus 8 -- 'bash profiles/scale_r1.sh 8'     (then 4, then 2)
N=${1:-8}
O=gpurun_out/scale_r1
mkdir -p $O
run() {  # tag, bench args
  local tag=$1; shift
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 \
    bench.py --gpus $N --steps 100 --warmup 5 --no-cpu "$@" > $O/$tag.json 2> $O/$tag.err
  echo "$tag rc=$? $(tail -1 $O/$tag.json | cut -c1-160)"
}
run 7b_q4_1_fused_n$N --model 7b --fmt q4_1
run 30b_q5_0_fused_n$N --model 30b --fmt q5_0
if [ "${2:-}" = nccl ]; then run 7b_q4_1_nccl_n$N --model 7b --fmt q4_1 --collective nccl; fi
if [ "$N" = 8 ]; then
  timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29534 \
    tests/multi_gpu_check.py > $O/check_n$N.log 2>&1; echo "check n=$N rc=$?"
fi
