#!/bin/bash
# Round 2, call ac: where did 1094 -> 1053 tok/s go?  The bench of three historical trees and of the current one, same box
set -u
O=gpurun_out/r2ac
mkdir -p $O
for c in 872b984 f95d6c9 5acf1df current; do
  if [ $c = current ]; then d=.; else d=build/wt_$c; fi
  extra="--no-cpu --no-e2e"
  (cd $d && timeout 300 python bench.py $extra $(python bench.py -h 2>/dev/null | grep -q -- --no-gemm && echo --no-gemm) $(python bench.py -h 2>/dev/null | grep -q -- --no-extras && echo --no-extras) --steps 200 > /tmp/b_$c.json 2>/tmp/b_$c.err)
  python -c "
import json
d=json.loads(open('/tmp/b_$c.json').read().strip().splitlines()[-1])
print('$c: value', round(d['value'],1), 'frac', round(d['roofline']['frac'],4), 'launches', d['config'].get('launches_per_step'))" | tee -a $O/bisect.txt
done
