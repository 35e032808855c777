#!/bin/bash
set -u
O=gpurun_out/r1j
mkdir -p $O
LAMM_SERVER_X_CTAS=8 timeout 300 python -m pytest tests -m gpu -q --timeout 200 -p no:cacheprovider -x -k "host_modes or operator" > $O/pytest_x8.log 2>&1; echo "pytest (x_ctas=8) rc=$? $(tail -1 $O/pytest_x8.log)"
grep -E "FAILED|ERROR" $O/pytest_x8.log | head
for t in 1 4 8 16 32; do
  LAMM_SERVER_X_CTAS=$t LAMM_SERVER_TRACE=1 timeout 90 python tools/trace_server.py >> $O/trace_server_x$t.txt 2>&1
  LAMM_SERVER_X_CTAS=$t LAMM_SERVER_TRACE=1 timeout 90 python tools/trace_server.py --m 4096 --k 11008 >> $O/trace_server_x$t.txt 2>&1
  grep -E "host wall|quantized|chunk" $O/trace_server_x$t.txt
  LAMM_SERVER_X_CTAS=$t timeout 200 python bench.py --steps 200 --warmup 5 --no-cpu > $O/bench_x$t.json 2> $O/bench_x$t.err; echo "bench x_ctas=$t rc=$?"
  python - $O/bench_x$t.json <<'PY'
import json, sys
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1], "value %.1f tok/s | e2e %.1f tok/s %.2f us/call" % (d["value"], d["e2e"]["value"], d["e2e"]["per_call_phases"]["total_per_call_us"]))
PY
done
