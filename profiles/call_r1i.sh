#!/bin/bash
set -u
O=gpurun_out/r1i
mkdir -p $O
LAMM_SERVER_TMA_X=1 timeout 300 python -m pytest tests -m gpu -q --timeout 200 -p no:cacheprovider -x -k "host_modes or operator" > $O/pytest_tma.log 2>&1; echo "pytest (tma_x) rc=$? $(tail -1 $O/pytest_tma.log)"
grep -E "FAILED|ERROR" $O/pytest_tma.log | head
for t in 0 1; do
  LAMM_SERVER_TMA_X=$t LAMM_SERVER_TRACE=1 timeout 90 python tools/trace_server.py >> $O/trace_server_tma$t.txt 2>&1
  LAMM_SERVER_TMA_X=$t LAMM_SERVER_TRACE=1 timeout 90 python tools/trace_server.py --m 4096 --k 11008 >> $O/trace_server_tma$t.txt 2>&1
  cat $O/trace_server_tma$t.txt
  LAMM_SERVER_TMA_X=$t timeout 200 python bench.py --steps 200 --warmup 5 --no-cpu > $O/bench_tma$t.json 2> $O/bench_tma$t.err; echo "bench tma_x=$t rc=$?"
  python - $O/bench_tma$t.json <<'PY'
import json, sys
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1], "value %.1f tok/s frac %.3f | e2e %.1f tok/s %.2f us/call" % (d["value"], d["roofline"]["frac"], d["e2e"]["value"], d["e2e"]["per_call_phases"]["total_per_call_us"]))
PY
done
