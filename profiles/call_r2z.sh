#!/bin/bash
# Round 2, call z: bench line with the multi-GPU workload keys at N = 1, GEMM suite re-check
set -u
O=gpurun_out/r2z
mkdir -p $O
timeout 900 python bench.py --no-cpu > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"; tail -3 $O/bench.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2z/bench.json').read().strip().splitlines()[-1])
print('value', round(d['value'],1), 'frac', round(d['roofline']['frac'],4))
print('gemm', [(c['fmt'], c['TOPS']) for c in d['gemm']['cases']])
print('extras', json.dumps(d.get('multi_gpu_workloads')))
PY
