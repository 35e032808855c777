#!/bin/bash
# Round 2, call s: SURVEY 8f-3 fusions (RMS-norm / SiLU-mul in front of the matmul) -- parity, then a decode bench re-check
set -u
O=gpurun_out/r2s
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_xop.py -m gpu -q -x --timeout 120 -p no:cacheprovider > $O/pytest_xop.log 2>&1
rc=$?; echo "xop tests rc=$rc $(tail -1 $O/pytest_xop.log)"; grep -E "^FAILED|^ERROR|Error|assert|err " $O/pytest_xop.log | head -20
if [ $rc -eq 124 ]; then exit 0; fi
timeout 300 python bench.py --no-cpu --no-e2e --no-gemm > $O/bench_quick.json 2> $O/bench_quick.err; echo "bench rc=$?"; tail -2 $O/bench_quick.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2s/bench_quick.json').read().strip().splitlines()[-1])
print('value', round(d['value'],1), 'frac', round(d['roofline']['frac'],4), 'sustained', d.get('sustained',{}).get('value'))
PY
timeout 200 python tools/xop_bench.py > $O/xop_bench.txt 2>&1; cat $O/xop_bench.txt
