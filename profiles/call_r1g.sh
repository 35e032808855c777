#!/bin/bash
set -u
O=gpurun_out/r1g
mkdir -p $O
for a in 0 1 2 4 8; do
  LAMM_DECODE_L2_AHEAD=$a timeout 200 python bench.py --steps 200 --warmup 5 --no-cpu > $O/bench_l2a$a.json 2> $O/bench_l2a$a.err; echo "bench l2_ahead=$a rc=$?"
  python - $O/bench_l2a$a.json <<'PY'
import json, sys
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1], "value %.1f tok/s frac %.3f | e2e %.1f tok/s" % (d["value"], d["roofline"]["frac"], d["e2e"]["value"]))
PY
done
LAMM_DECODE_L2_AHEAD=2 timeout 600 python -m pytest tests -m gpu -q --timeout 200 -p no:cacheprovider -x -k "decode or kquant or llama7b or golden" > $O/pytest_gpu.log 2>&1; echo "pytest (l2_ahead=2) rc=$? $(tail -1 $O/pytest_gpu.log)"
for a in 2 4; do
LAMM_DECODE_L2_AHEAD=$a timeout 300 python bench.py --suite gemv --fused --fmts q4_1,q8_0,q4_k --shapes 4096x4096,11008x4096,4096x11008,32000x4096,17920x6656 > $O/sweep_decode_l2a$a.jsonl 2>&1; echo "sweep rc=$?"
cut -c1-150 $O/sweep_decode_l2a$a.jsonl
done
