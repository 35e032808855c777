#!/bin/bash
# Round 2, call bq: job-tag restart before the 32-bit wrap -- operator tests incl. the restart test, e2e line
set -u
O=gpurun_out/r2bq
mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_ggml_hook.py -m gpu -q -x --timeout 300 -p no:cacheprovider -k "operator or hook or ggml" > $O/pytest_operator.log 2>&1
echo "operator + hook tests rc=$? $(tail -1 $O/pytest_operator.log)"; grep -E "^FAILED|^ERROR|assert " $O/pytest_operator.log | head -5
timeout 300 python bench.py --no-gemm --no-extras --sustain-s 0 > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"
python - <<PY
import json
d=json.load(open('$O/bench.json'))
print('value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'cpu', d['cpu_baseline']['value'], d['cpu_baseline']['per_shape_us'])
PY
