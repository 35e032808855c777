#!/bin/bash
# Round 2, call bm: resident server L2 prefetch (bit 0: own chunk during the X pull, bit 1: next sibling job's chunk) -- e2e A/B
set -u
O=gpurun_out/r2bm
mkdir -p $O
for pf in 0 1 2 3 0 1 2; do
  LAMM_SERVER_L2_PREFETCH=$pf timeout 300 python bench.py --no-gemm --no-extras --sustain-s 0 --no-cpu --steps 5 --warmup 3 > $O/bench_pf$pf.json 2> $O/bench_pf$pf.err
  python - <<PY | tee -a $O/ab.txt
import json
d=json.load(open('$O/bench_pf$pf.json'))
print('prefetch=$pf e2e', round(d['e2e']['value'],1), 'wait', d['e2e']['per_call_phases']['wait_us'], 'total', d['e2e']['per_call_phases']['total_per_call_us'])
PY
done
