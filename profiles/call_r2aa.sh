#!/bin/bash
# Round 2, call aa (2 GPUs): the multi-GPU pytest (row-sharded decode / prefill / sharded GGUF load) and the bench at N = 2
set -u
O=gpurun_out/r2aa
mkdir -p $O
nvidia-smi -L | head -4
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -q -x --timeout 800 -p no:cacheprovider > $O/pytest_multi.log 2>&1
echo "multi tests rc=$? $(tail -1 $O/pytest_multi.log)"; grep -E "FAIL|Error|assert" $O/pytest_multi.log | head -10
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29711 tests/multi_gpu_check.py > $O/check_n2.log 2>&1
echo "check rc=$?"; grep -c " ok" $O/check_n2.log; grep -E "gguf|FAIL|all ok" $O/check_n2.log | head -24
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29712 bench.py --gpus 2 --steps 200 --warmup 5 > $O/bench_n2.json 2> $O/bench_n2.err
echo "bench n2 rc=$?"; tail -3 $O/bench_n2.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2aa/bench_n2.json').read().strip().splitlines()[-1])
print('value', round(d['value'],1), 'n_gpus', d['n_gpus'], 'e2e', d['e2e'].get('value'))
print('extras', json.dumps(d.get('multi_gpu_workloads')))
PY
