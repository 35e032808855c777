#!/bin/bash
# Round 2, call bi: resident server with a mailbox ring + sibling speculation -- hang guard, host-path parity, e2e rate
set -u
O=gpurun_out/r2bi
mkdir -p $O
timeout 120 python tools/sanitize_small.py 3,10 > $O/small.txt 2>&1; rc=$?; echo "small rc=$rc $(tail -1 $O/small.txt)"
if [ $rc -ne 0 ]; then tail -20 $O/small.txt; exit 0; fi
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x --timeout 300 -p no:cacheprovider -k "operator" > $O/pytest_operator.log 2>&1
echo "operator tests rc=$? $(tail -1 $O/pytest_operator.log)"; grep -E "^FAILED|^ERROR|Error|assert " $O/pytest_operator.log | head -10
timeout 600 python bench.py --no-gemm --no-extras --sustain-s 0 > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2bi/bench.json'))
print('value', d['value'], 'e2e', d['e2e'])
PY
LAMM_SERVER_SPEC=0 timeout 600 python bench.py --no-gemm --no-extras --sustain-s 0 --no-cpu > $O/bench_nospec.json 2> $O/bench_nospec.err; echo "bench (no speculation) rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2bi/bench_nospec.json'))
print('nospec e2e', d['e2e'])
PY
