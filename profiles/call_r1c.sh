#!/bin/bash
# Round-1 second session, measurement call C: K-quant units in the single-launch decode kernel.
set -u
O=gpurun_out/r1c
mkdir -p $O
timeout 600 python -m pytest tests -m gpu -q --timeout 200 -p no:cacheprovider > $O/pytest_gpu.log 2>&1; echo "pytest rc=$? $(tail -1 $O/pytest_gpu.log)"
grep -E "FAILED|ERROR" $O/pytest_gpu.log | head -30
timeout 300 python bench.py --suite gemv --fused --fmts q2_k,q3_k,q4_k,q5_k,q6_k --shapes 4096x4096,11008x4096,4096x11008,32000x4096,17920x6656 > $O/sweep_kquant_decode.jsonl 2>&1; echo "kquant decode sweep rc=$?"
cut -c1-150 $O/sweep_kquant_decode.jsonl
timeout 200 python bench.py --steps 100 --warmup 5 --no-cpu --chain ll > $O/bench_chain_ll.json 2> $O/bench_chain_ll.err; echo "chain ll rc=$?"
LAMM_DECODE_CTAS_PER_SM=2 timeout 200 python bench.py --steps 100 --warmup 5 --no-cpu > $O/bench_2cta.json 2> $O/bench_2cta.err; echo "2 CTA/SM rc=$?"
for f in $O/bench_*.json; do python - "$f" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], "value %.1f tok/s frac %.3f launches/step %s | e2e %.1f" % (d["value"], d["roofline"]["frac"], d["config"].get("launches_per_step"), d["e2e"]["value"]))
except Exception as ex:
    print(sys.argv[1], "unreadable:", ex)
PY
done
