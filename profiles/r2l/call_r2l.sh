#!/bin/bash
# Round 2, call l: Q2_K drain conversion variants (LAMM_Q2K_CVT 0 = I2FP, 1 = magic, 2 = half) -- temporary switch
set -u
O=gpurun_out/r2l
mkdir -p $O
for v in 0 1 2; do
  LAMM_Q2K_CVT=$v timeout 60 python tools/sanitize_small.py 10 > $O/small_cvt$v.log 2>&1; echo "cvt $v small rc=$? $(tail -1 $O/small_cvt$v.log)"
  LAMM_Q2K_CVT=$v timeout 100 python bench.py --suite gemm --n 512 --fmts q2_k --shapes 4096x4096,11008x4096,4096x11008 > $O/n512_cvt$v.jsonl 2>&1
  LAMM_Q2K_CVT=$v timeout 60 python bench.py --suite gemm --n 2048 --fmts q2_k --shapes 5120x5120,13824x5120,5120x13824 > $O/n2048_cvt$v.jsonl 2>&1
done
python - <<'P'
import json
for v in (0, 1, 2):
    out = []
    for name in ("n512", "n2048"):
        for line in open("gpurun_out/r2l/%s_cvt%d.jsonl" % (name, v)):
            try: d = json.loads(line)
            except ValueError: continue
            if "fmt" in d: out.append("%sx%sx%s %.0f" % (d.get("M"), d.get("K"), d.get("N"), d.get("TOPS")))
    print("cvt", v, " | ".join(out))
P
