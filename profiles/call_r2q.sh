#!/bin/bash
# Round 2, call q: persistent chain with 32 warps per CTA
set -u
O=gpurun_out/r2q
mkdir -p $O
timeout 300 python -m pytest tests/test_gpu_chain.py -m gpu -q -x --timeout 120 -p no:cacheprovider > $O/pytest_chain.log 2>&1
rc=$?; echo "chain tests rc=$rc $(tail -1 $O/pytest_chain.log)"; grep -E "^FAILED|^ERROR|Error|assert" $O/pytest_chain.log | head -10
if [ $rc -ne 0 ]; then exit 0; fi
timeout 120 python tools/trace_chain.py > $O/trace_chain.txt 2>&1; cat $O/trace_chain.txt
for mode in persistent plain; do
  timeout 200 python bench.py --chain $mode --no-cpu --no-e2e --no-gemm > $O/bench_$mode.json 2> $O/bench_$mode.err; echo "bench $mode rc=$?"; tail -2 $O/bench_$mode.err
  python -c "
import json
d=json.loads(open('$O/bench_$mode.json').read().strip().splitlines()[-1])
print('$mode', 'value', round(d['value'],1), 'frac', round(d['roofline']['frac'],4), 'sustained', round(d['sustained']['value'],1))"
done
