#!/bin/bash
set -u
O=gpurun_out/r1k
mkdir -p $O
LAMM_SERVER_IO_CTA=1 timeout 300 python -m pytest tests -m gpu -q --timeout 200 -p no:cacheprovider -x -k "host_modes or operator" > $O/pytest_io.log 2>&1; echo "pytest (io_cta) rc=$? $(tail -1 $O/pytest_io.log)"
grep -E "FAILED|ERROR" $O/pytest_io.log | head
for t in 0 1; do
  LAMM_SERVER_IO_CTA=$t LAMM_SERVER_TRACE=1 timeout 90 python tools/trace_server.py >> $O/trace_server_io$t.txt 2>&1
  LAMM_SERVER_IO_CTA=$t LAMM_SERVER_TRACE=1 timeout 90 python tools/trace_server.py --m 4096 --k 11008 >> $O/trace_server_io$t.txt 2>&1
  cat $O/trace_server_io$t.txt
  for rep in 1 2; do
  LAMM_SERVER_IO_CTA=$t timeout 200 python bench.py --steps 200 --warmup 5 --no-cpu > $O/bench_io${t}_$rep.json 2> $O/bench_io$t.err; echo "bench io_cta=$t rc=$?"
  python - $O/bench_io${t}_$rep.json <<'PY'
import json, sys
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1], "value %.1f tok/s | e2e %.1f tok/s %.2f us/call" % (d["value"], d["e2e"]["value"], d["e2e"]["per_call_phases"]["total_per_call_us"]))
PY
  done
done
