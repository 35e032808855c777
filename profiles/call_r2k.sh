#!/bin/bash
# Round 2, call k: Q2_K kind::i8 kernel, unpack without local-memory indexing, cheap copy addressing
set -u
O=gpurun_out/r2k
mkdir -p $O
timeout 90 python tools/sanitize_small.py 10 > $O/small_q2k.log 2>&1; rc=$?; echo "small q2_k rc=$rc $(tail -1 $O/small_q2k.log)"; grep FAIL $O/small_q2k.log | head
if [ $rc -eq 124 ]; then echo "HANG: stopping"; exit 0; fi
timeout 300 python -m pytest tests/test_gpu_gemm.py tests/test_gpu_fullsize.py -m gpu -q -x --timeout 120 -p no:cacheprovider -k "q2k or 10 or kquants" > $O/pytest.log 2>&1
echo "pytest rc=$? $(tail -1 $O/pytest.log)"; grep -E "^FAILED|^ERROR" $O/pytest.log | head
timeout 100 python bench.py --suite gemm --n 512 --fmts q2_k --shapes 4096x4096,11008x4096,4096x11008 > $O/sweep_q2k_n512.jsonl 2>&1; echo "sweep n512 rc=$?"
timeout 60 python bench.py --suite gemm --n 2048 --fmts q2_k --shapes 5120x5120,13824x5120,5120x13824 > $O/sweep_q2k_n2048.jsonl 2>&1; echo "sweep 13b rc=$?"
python - <<'P'
import json
for name in ("q2k_n512", "q2k_n2048"):
    for line in open("gpurun_out/r2k/sweep_%s.jsonl" % name):
        try: d = json.loads(line)
        except ValueError: continue
        if "fmt" in d: print(d["fmt"], d.get("M"), d.get("K"), d.get("N"), d.get("us"), d.get("TOPS"))
P
timeout 60 python tools/gemm_one.py q2_k 4096 4096 512 0 > $O/plain_q2k.log 2>&1 && \
timeout 300 ncu --set full --clock-control none --import-source on -k regex:gemm_q2k -s 3 -c 1 -f -o $O/gemm_q2k_i8 \
   python tools/gemm_one.py q2_k 4096 4096 512 0 > $O/ncu_q2k.log 2>&1
echo "q2k capture rc=$?"; cat $O/plain_q2k.log
