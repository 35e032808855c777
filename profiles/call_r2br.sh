#!/bin/bash
# Round 2, call br: server X pull with every load of the row in flight at once (one PCIe round trip for K = 11008) -- e2e A/B on one box
set -u
O=gpurun_out/r2br
mkdir -p $O
for v in new head new head; do
  if [ $v = head ]; then export LAMM_LIB=$PWD/build_variants/liblammcu_head.so; else unset LAMM_LIB; fi
  timeout 300 python bench.py --no-gemm --no-extras --sustain-s 0 --no-cpu --steps 5 --warmup 3 > $O/bench_$v.json 2> $O/bench_$v.err
  python - <<PY | tee -a $O/ab.txt
import json
d=json.load(open('$O/bench_$v.json'))
print('$v e2e', round(d['e2e']['value'],1), 'wait', d['e2e']['per_call_phases']['wait_us'], 'total', d['e2e']['per_call_phases']['total_per_call_us'])
PY
done
