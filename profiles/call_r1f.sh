#!/bin/bash
set -u
O=gpurun_out/${1:-r1f}
mkdir -p $O
timeout 600 python -m pytest tests -m gpu -q --timeout 200 -p no:cacheprovider -x > $O/pytest_gpu.log 2>&1; echo "pytest rc=$? $(tail -1 $O/pytest_gpu.log)"
grep -E "FAILED|ERROR" $O/pytest_gpu.log | head -30
timeout 200 python bench.py --steps 200 --warmup 5 --no-cpu > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"
python - $O <<'PY'
import json, sys
d = json.loads(open(sys.argv[1] + "/bench.json").read().strip().splitlines()[-1])
print("value %.1f tok/s frac %.3f | e2e %.1f tok/s %.2f us/call" % (d["value"], d["roofline"]["frac"], d["e2e"]["value"], d["e2e"]["per_call_phases"]["total_per_call_us"]))
PY
timeout 300 python bench.py --suite gemv --fused --fmts q4_0,q4_1,q5_1,q8_0,q2_k --shapes 4096x4096,11008x4096,4096x11008,32000x4096,17920x6656 > $O/sweep_decode.jsonl 2>&1; echo "sweep rc=$?"
cut -c1-150 $O/sweep_decode.jsonl
