#!/bin/bash
# Round 2, call bn (8 GPUs, final kernels): the bench line at N = 8 with the multi-GPU workload keys; multi-GPU parity script
set -u
O=gpurun_out/r2bn
mkdir -p $O
nvidia-smi -L | wc -l
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29721 bench.py --gpus 8 --steps 50 --warmup 3 > $O/bench_n8.json 2> $O/bench_n8.err
echo "bench n8 rc=$?"; tail -3 $O/bench_n8.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2bn/bench_n8.json').read().strip().splitlines()[-1])
print('value', round(d['value'],1), 'n_gpus', d['n_gpus'])
print('extras', json.dumps(d.get('multi_gpu_workloads')))
PY
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29722 tests/multi_gpu_check.py > $O/check_n8.log 2>&1
echo "check rc=$?"; grep -c " ok" $O/check_n8.log; grep -E "FAIL|all ok" $O/check_n8.log | head
