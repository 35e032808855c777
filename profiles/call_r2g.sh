#!/bin/bash
# Round 2, call g (after the machine was replaced: re-verifies everything rebuilt from f95d6c9): hang guard on the
# Q2_K kind::i8 kernel, the whole GPU suite, the tensor-pipe peaks, Q2_K prefill sweeps.
set -u
O=gpurun_out/r2g
mkdir -p $O
timeout 90 python tools/sanitize_small.py 10 > $O/small_q2k.log 2>&1; rc=$?; echo "small q2_k rc=$rc $(tail -1 $O/small_q2k.log)"; grep FAIL $O/small_q2k.log | head
if [ $rc -eq 124 ]; then echo "HANG: stopping"; exit 0; fi
timeout 900 python -m pytest tests -m gpu -q --timeout 300 -p no:cacheprovider > $O/pytest_gpu.log 2>&1
echo "pytest rc=$? $(tail -1 $O/pytest_gpu.log)"; grep -E "^FAILED|^ERROR" $O/pytest_gpu.log | head -30
timeout 120 scratch/i8peak > $O/i8peak.txt 2>&1; echo "i8peak rc=$?"; cat $O/i8peak.txt
timeout 100 python bench.py --suite gemm --n 512 --fmts q2_k --shapes 4096x4096,11008x4096,4096x11008 > $O/sweep_q2k_n512.jsonl 2>&1; echo "sweep n512 rc=$?"
timeout 60 python bench.py --suite gemm --n 2048 --fmts q2_k,q4_1 --shapes 5120x5120,13824x5120,5120x13824 > $O/sweep_13b_n2048.jsonl 2>&1; echo "sweep 13b rc=$?"
python - <<'P'
import json
for name in ("q2k_n512", "13b_n2048"):
    for line in open("gpurun_out/r2g/sweep_%s.jsonl" % name):
        try: d = json.loads(line)
        except ValueError: continue
        if "fmt" in d: print(d["fmt"], d.get("M"), d.get("K"), d.get("N"), d.get("us"), d.get("TOPS"))
P
