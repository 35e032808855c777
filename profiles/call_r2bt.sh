#!/bin/bash
# Round 2, call bt: server CTAs stash the next job seen during the current poll (no L2 round trip for a follower) -- tests, e2e A/B on one box
set -u
O=gpurun_out/r2bt
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_ggml_hook.py -m gpu -q -x --timeout 300 -p no:cacheprovider -k "operator or hook or ggml" > $O/pytest_operator.log 2>&1
echo "operator + hook tests rc=$? $(tail -1 $O/pytest_operator.log)"; grep -E "^FAILED|^ERROR|assert " $O/pytest_operator.log | head -5
for v in new head new head; do
  if [ $v = head ]; then export LAMM_LIB=$PWD/build_variants/liblammcu_head.so; else unset LAMM_LIB; fi
  timeout 300 python bench.py --no-gemm --no-extras --sustain-s 0 --no-cpu --steps 5 --warmup 3 > $O/bench_$v.json 2> $O/bench_$v.err
  python - <<PY | tee -a $O/ab.txt
import json
d=json.load(open('$O/bench_$v.json'))
print('$v e2e', round(d['e2e']['value'],1), 'wait', d['e2e']['per_call_phases']['wait_us'], 'total', d['e2e']['per_call_phases']['total_per_call_us'])
PY
done
unset LAMM_LIB
timeout 100 python tools/trace_server.py --siblings 3 2>&1 | tail -3 | tee $O/trace_siblings.txt
