#!/bin/bash
set -u
O=gpurun_out/r1e
mkdir -p $O
timeout 600 python -m pytest tests -m gpu -q --timeout 200 -p no:cacheprovider -x > $O/pytest_gpu.log 2>&1; echo "pytest rc=$? $(tail -1 $O/pytest_gpu.log)"
grep -E "FAILED|ERROR" $O/pytest_gpu.log | head -30
timeout 200 python bench.py --steps 200 --warmup 5 --no-cpu > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r1e/bench.json").read().strip().splitlines()[-1])
print("value %.1f tok/s frac %.3f | e2e %.1f tok/s %.2f us/call" % (d["value"], d["roofline"]["frac"], d["e2e"]["value"], d["e2e"]["per_call_phases"]["total_per_call_us"]))
PY
timeout 90 python tools/trace_decode.py > $O/trace_decode.txt 2>&1; echo "decode trace rc=$?"
cat $O/trace_decode.txt
