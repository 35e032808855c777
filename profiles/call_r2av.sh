#!/bin/bash
# Round 2, call av: F16 streaming kernel with fp16 activations in shared memory -- parity, decode sweep
set -u
O=gpurun_out/r2av
mkdir -p $O
timeout 300 python -m pytest tests/test_gpu_f16.py tests/test_gpu_xop.py -m gpu -q -x --timeout 90 -p no:cacheprovider -k "f16 or F16 or 1-" > $O/pytest.log 2>&1; echo "pytest rc=$? $(tail -1 $O/pytest.log)"
timeout 300 python -m pytest tests/test_gpu_f16.py -m gpu -q --timeout 90 -p no:cacheprovider > $O/pytest_f16.log 2>&1; echo "f16 rc=$? $(tail -1 $O/pytest_f16.log)"
timeout 300 python bench.py --suite gemv --fused --fmts f16,f32 > $O/sweep_f16.jsonl 2>&1
python - <<'PY'
import json
for l in open('gpurun_out/r2av/sweep_f16.jsonl'):
    try: d=json.loads(l)
    except: continue
    print(d['fmt'], d['M'], d['K'], round(d['us'],1), 'us', round(100*d['frac_hbm']), '%')
PY
