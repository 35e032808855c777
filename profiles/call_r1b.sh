#!/bin/bash
# Round-1 second session, measurement call B (run on the GPU box from the repo root):
#   gpurun --timeout 840 -- 'bash profiles/call_r1b.sh'
# parity of everything (incl. the K-quants and host mode 2), host-path variants of the e2e leg,
# K-quant sweeps, host-latency floor and in-kernel time lines.  Every step has its own timeout.
set -u
O=gpurun_out/r1b
mkdir -p $O
timeout 500 python -m pytest tests -m gpu -q --timeout 200 -p no:cacheprovider > $O/pytest_gpu.log 2>&1; echo "pytest rc=$? $(tail -1 $O/pytest_gpu.log)"
nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o $O/hostlat tools/hostlat.cu 2>/dev/null && timeout 60 $O/hostlat > $O/hostlat.txt 2>&1; echo "hostlat rc=$?"; cat $O/hostlat.txt
timeout 300 python bench.py --steps 200 --warmup 5 --host-mode 1 > $O/bench_mode1.json 2> $O/bench_mode1.err; echo "bench mode1 rc=$?"
LAMM_SERVER_SENTINEL=1 timeout 240 python bench.py --steps 100 --warmup 5 --no-cpu --host-mode 1 > $O/bench_mode1_sentinel.json 2> $O/bench_mode1_sentinel.err; echo "bench mode1+sentinel rc=$?"
timeout 240 python bench.py --steps 100 --warmup 5 --no-cpu --host-mode 2 > $O/bench_mode2.json 2> $O/bench_mode2.err; echo "bench mode2 rc=$?"
LAMM_SERVER_SENTINEL=1 timeout 240 python bench.py --steps 100 --warmup 5 --no-cpu --host-mode 2 > $O/bench_mode2_sentinel.json 2> $O/bench_mode2_sentinel.err; echo "bench mode2+sentinel rc=$?"
LAMM_BENCH_PYLOOP=1 timeout 240 python bench.py --steps 100 --warmup 5 --no-cpu --host-mode 1 > $O/bench_mode1_pyloop.json 2> $O/bench_mode1_pyloop.err; echo "bench mode1 python loop rc=$?"
for f in $O/bench_mode*.json; do python - "$f" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    e = d["e2e"]
    print(sys.argv[1], "value %.1f tok/s frac %.3f | e2e %.1f tok/s, %.2f us/call, mode %s sentinel %s caller %s" % (
        d["value"], d["roofline"]["frac"], e["value"], e["per_call_phases"]["total_per_call_us"], e.get("host_mode"),
        e.get("server_sentinel"), e.get("caller")))
except Exception as ex:
    print(sys.argv[1], "unreadable:", ex)
PY
done
timeout 200 python bench.py --suite gemv --fused --fmts q3_k,q4_k,q5_k,q6_k --shapes 4096x4096,11008x4096,4096x11008,32000x4096 > $O/sweep_kquant_gemv.jsonl 2>&1; echo "kquant gemv sweep rc=$?"
timeout 200 python bench.py --suite gemm --n 512 --fmts q3_k,q4_k,q5_k,q6_k --shapes 4096x4096,11008x4096 > $O/sweep_kquant_gemm_n512.jsonl 2>&1; echo "kquant gemm sweep rc=$?"
for m in 1 2; do for s in 0 1; do
  LAMM_SERVER_TRACE=1 timeout 90 python tools/trace_server.py --mode $m --sentinel $s >> $O/trace_server.txt 2>&1
  LAMM_SERVER_TRACE=1 timeout 90 python tools/trace_server.py --mode $m --sentinel $s --m 4096 --k 11008 >> $O/trace_server.txt 2>&1
done; done; echo "server traces rc=$?"
timeout 90 python tools/trace_decode.py > $O/trace_decode.txt 2>&1; echo "decode trace rc=$?"
timeout 90 python tools/trace_decode.py --m 11008 --k 11008 --chain 6 >> $O/trace_decode.txt 2>&1
tail -40 $O/trace_server.txt
tail -20 $O/trace_decode.txt
cat $O/sweep_kquant_gemv.jsonl | cut -c1-150
cat $O/sweep_kquant_gemm_n512.jsonl | cut -c1-170
grep -E "FAILED|ERROR|passed|failed" $O/pytest_gpu.log | tail -30
