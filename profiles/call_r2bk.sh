#!/bin/bash
# Round 2, call bk: write-combined mailbox + activation row for the resident server (LAMM_SERVER_WC): e2e with / without, operator tests with
set -u
O=gpurun_out/r2bk
mkdir -p $O
for wc in 0 1 0 1; do
  LAMM_SERVER_WC=$wc timeout 300 python bench.py --no-gemm --no-extras --sustain-s 0 --no-cpu --steps 5 --warmup 3 > $O/bench_wc$wc.json 2> $O/bench_wc$wc.err
  python - <<PY
import json
d=json.load(open('$O/bench_wc$wc.json'))
print('WC=$wc e2e', round(d['e2e']['value'],1), d['e2e']['per_call_phases'])
PY
done
LAMM_SERVER_WC=1 timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -x --timeout 300 -p no:cacheprovider -k "operator" > $O/pytest_operator_wc.log 2>&1
echo "operator tests (WC) rc=$? $(tail -1 $O/pytest_operator_wc.log)"
