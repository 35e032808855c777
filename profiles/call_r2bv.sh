#!/bin/bash
# Round 2, call bv: the sibling bookkeeping moved into lamm_sibling.h (CPU-tested) -- operator + hook tests and the e2e line on hardware
set -u
O=gpurun_out/r2bv
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_ggml_hook.py -m gpu -q -x --timeout 300 -p no:cacheprovider -k "operator or hook or ggml" > $O/pytest_operator.log 2>&1
echo "operator + hook tests rc=$? $(tail -1 $O/pytest_operator.log)"; grep -E "^FAILED|^ERROR|assert " $O/pytest_operator.log | head -5
timeout 300 python bench.py --no-gemm --no-extras --sustain-s 0 --no-cpu --steps 5 --warmup 3 > $O/bench.json 2> $O/bench.err
python - <<PY
import json
d=json.load(open('$O/bench.json'))
print('e2e', round(d['e2e']['value'],1), d['e2e']['per_call_phases'])
PY
