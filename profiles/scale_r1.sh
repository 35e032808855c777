#!/bin/bash
# Round-1 multi-GPU scaling capture: gpurun --gpus 8 -- 'bash profiles/scale_r1.sh 8'
NMAX=${1:-8}
O=gpurun_out/scale_r1
mkdir -p $O
run() {  # n, extra args, tag
  local n=$1; shift; local tag=$1; shift
  if [ "$n" = 1 ]; then
    timeout 300 python bench.py --gpus 1 --steps 100 --warmup 5 --no-cpu "$@" > $O/$tag.json 2> $O/$tag.err
  else
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29533 \
      bench.py --gpus $n --steps 100 --warmup 5 --no-cpu "$@" > $O/$tag.json 2> $O/$tag.err
  fi
  echo "$tag rc=$? $(tail -1 $O/$tag.json | cut -c1-160)"
}
for n in 1 2 4 8; do
  [ $n -le $NMAX ] || continue
  run $n 7b_q4_1_fused_n$n --model 7b --fmt q4_1
  run $n 30b_q5_0_fused_n$n --model 30b --fmt q5_0
  if [ $n -gt 1 ]; then run $n 7b_q4_1_nccl_n$n --model 7b --fmt q4_1 --collective nccl; fi
done
if [ $NMAX -ge 2 ]; then
  timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NMAX --master-addr 127.0.0.1 --master-port 29534 \
    tests/multi_gpu_check.py > $O/check_n$NMAX.log 2>&1; echo "check n=$NMAX rc=$?"
fi
