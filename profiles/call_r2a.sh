#!/bin/bash
# First GPU call of round 2 (prepared at the end of round 1, when the GPU budget was spent):
#   gpurun --timeout 420 -- 'bash profiles/call_r2a.sh'        (about 2.5 GPU-minutes)
# 1. the regular GPU suite (must stay green), 2. the opt-in checks of everything written without a GPU
# (LAMM_DECODE_RING, lamm_mul_mat_id, GGUF rows -> device), 3. the ring-depth sweep that decides whether
# 3 / 4 slots become a per-format default, 4. the chain bench with each depth.
set -u
O=gpurun_out/r2a
mkdir -p $O
timeout 200 python -m pytest tests -m gpu -q --timeout 200 -p no:cacheprovider -x > $O/pytest_gpu.log 2>&1
echo "pytest rc=$? $(tail -1 $O/pytest_gpu.log)"
LAMM_TEST_EXPERIMENTAL=1 timeout 300 python -m pytest tests/test_gpu_experimental.py -m gpu -q --timeout 280 -p no:cacheprovider \
  > $O/pytest_experimental.log 2>&1
echo "experimental rc=$? $(tail -1 $O/pytest_experimental.log)"
grep -E "FAILED|ERROR|passed|failed" $O/pytest_experimental.log | head -20
for r in 2 3 4; do
  LAMM_DECODE_RING=$r timeout 150 python bench.py --suite gemv --fused --fmts q2_k,q3_k,q4_0,q4_1,q4_k,q8_0 \
    --shapes 4096x4096,11008x4096,4096x11008,32000x4096 > $O/sweep_ring$r.jsonl 2>&1
  echo "ring $r sweep rc=$?"
done
python - <<'P'
import json
rows = {}
for r in (2, 3, 4):
    for line in open("gpurun_out/r2a/sweep_ring%d.jsonl" % r):
        try:
            d = json.loads(line)
        except ValueError:
            continue
        if "fmt" in d:
            rows.setdefault((d["fmt"], d.get("M"), d.get("K")), {})[r] = (d.get("us"), d.get("frac_hbm"))
for key in sorted(rows):
    print(key, " ".join("ring%d=%s" % (r, rows[key].get(r)) for r in (2, 3, 4)))
P
LAMM_DECODE_PAIR=1 timeout 100 python bench.py --suite gemv --fused --fmts q4_0,q4_1 \
  --shapes 4096x4096,11008x4096,4096x11008,32000x4096 > $O/sweep_pair.jsonl 2>&1; echo "pair sweep rc=$?"; cut -c1-120 $O/sweep_pair.jsonl
LAMM_DECODE_PAIR=1 timeout 60 python bench.py --no-cpu --no-e2e > $O/bench_pair.json 2> $O/bench_pair.err
python -c "import json;d=json.loads(open('$O/bench_pair.json').read().strip().splitlines()[-1]);print('pair', d['value'], d['roofline']['frac'])"
for r in 2 3 4; do
  LAMM_DECODE_RING=$r timeout 60 python bench.py --no-cpu --no-e2e > $O/bench_ring$r.json 2> $O/bench_ring$r.err
  python -c "import json;d=json.loads(open('$O/bench_ring$r.json').read().strip().splitlines()[-1]);print('ring $r', d['value'], d['roofline']['frac'])"
done
