#!/bin/bash
set -u
O=gpurun_out/r1h
mkdir -p $O
timeout 600 python -m pytest tests -m gpu -q --timeout 200 -p no:cacheprovider -x > $O/pytest_gpu.log 2>&1; echo "pytest rc=$? $(tail -1 $O/pytest_gpu.log)"
grep -E "FAILED|ERROR" $O/pytest_gpu.log | head
timeout 300 python bench.py --suite gemv --fused --fmts q8_0,q6_k --shapes 4096x4096,11008x4096,4096x11008,32000x4096,5120x5120,13824x5120,6656x6656,17920x6656,6656x17920 > $O/sweep_decode_q8.jsonl 2>&1; echo "sweep rc=$?"
cut -c1-150 $O/sweep_decode_q8.jsonl
for f in q4_1 q8_0; do
timeout 200 python bench.py --steps 200 --warmup 5 --no-cpu --fmt $f > $O/bench_$f.json 2> $O/bench_$f.err; echo "bench $f rc=$?"
python - $O/bench_$f.json <<'PY'
import json, sys
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1], "value %.1f tok/s frac %.3f | e2e %.1f tok/s" % (d["value"], d["roofline"]["frac"], d["e2e"]["value"]))
PY
done
