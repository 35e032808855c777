#!/bin/bash
# Round 2, call ab: does the nvidia-smi clock sampler perturb the timed region?  bench with / without it, alternating
set -u
O=gpurun_out/r2ab
mkdir -p $O
for i in 1 2 3; do
  for s in on off; do
    if [ $s = off ]; then export LAMM_BENCH_NO_SAMPLER=1; else unset LAMM_BENCH_NO_SAMPLER; fi
    timeout 300 python bench.py --no-cpu --no-e2e --no-gemm --no-extras --sustain-s 0 > $O/b_${s}_$i.json 2>/dev/null
    python -c "
import json
d=json.loads(open('$O/b_${s}_$i.json').read().strip().splitlines()[-1])
print('sampler $s run $i: value', round(d['value'],1), 'frac', round(d['roofline']['frac'],4))" | tee -a $O/sampler.txt
  done
done
