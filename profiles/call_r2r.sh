#!/bin/bash
# Round 2, call r: F16 weights -- hang guard, parity, a prefill rate, then the whole GPU suite and the bench
set -u
O=gpurun_out/r2r
mkdir -p $O
timeout 400 python -m pytest tests/test_gpu_f16.py -m gpu -q -x --timeout 90 -p no:cacheprovider > $O/pytest_f16.log 2>&1
rc=$?; echo "f16 tests rc=$rc $(tail -1 $O/pytest_f16.log)"; grep -E "^FAILED|^ERROR|Error|assert|err " $O/pytest_f16.log | head -20
if [ $rc -eq 124 ]; then exit 0; fi
for s in "4096 4096 512" "11008 4096 2048" "13824 5120 2048"; do
  timeout 120 python tools/gemm_one.py f16 $s 2>&1 | tail -1 | tee -a $O/gemm_f16.txt
done
timeout 1500 python -m pytest tests -m gpu -q --timeout 300 -p no:cacheprovider > $O/pytest_gpu.log 2>&1
echo "gpu suite rc=$? $(tail -1 $O/pytest_gpu.log)"; grep -E "^FAILED|^ERROR" $O/pytest_gpu.log | head -20
timeout 600 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"; tail -3 $O/bench.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2r/bench.json').read().strip().splitlines()[-1])
print('value', round(d['value'],1), 'frac', round(d['roofline']['frac'],4), 'e2e', d.get('e2e'), 'clocks', d.get('clocks'))
print('gemm', json.dumps(d.get('gemm'))[:600])
PY
