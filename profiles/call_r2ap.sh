#!/bin/bash
# Round 2, call ap: decode variants -- va = fused-op code in its own instantiations (XM 4); vb = va + round-1 template structure
set -u
O=gpurun_out/r2ar
mkdir -p $O
for v in vb vi vf vb vi vf; do
  if [ $v = current ]; then L=""; else L="$PWD/build_variants/liblammcu_$v.so"; fi
  LAMM_LIB=$L timeout 300 python bench.py --no-cpu --no-e2e --no-gemm --no-extras --sustain-s 0 > $O/b_$v.json 2>$O/b_$v.err; tail -1 $O/b_$v.err
  python -c "
import json
d=json.loads(open('$O/b_$v.json').read().strip().splitlines()[-1])
print('$v: value', round(d['value'],1), 'frac', round(d['roofline']['frac'],4))" | tee -a $O/result.txt
done
