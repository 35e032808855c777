#!/bin/bash
# Round 2, call as: decode kernel with the fused-op code in its own instantiations (XM 4) and the disabled L2-lookahead
# branch back in the refill path: parity of everything that touches the decode kernel, then the bench
set -u
O=gpurun_out/r2as
mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_xop.py tests/test_gpu_kquants.py tests/test_gpu_chain.py tests/test_gpu_fullsize.py tests/test_zz_no_alloc.py -m gpu -q -x --timeout 200 -p no:cacheprovider > $O/pytest.log 2>&1; echo "pytest rc=$? $(tail -1 $O/pytest.log)"
timeout 300 python bench.py --no-cpu --no-e2e --no-gemm --no-extras > $O/b.json 2>$O/b.err; tail -1 $O/b.err
python -c "
import json
d=json.loads(open('$O/b.json').read().strip().splitlines()[-1])
print('value', round(d['value'],1), 'frac', round(d['roofline']['frac'],4), 'sustained', round(d['sustained']['value'],1))" | tee -a $O/result.txt
LAMM_DECODE_L2_AHEAD=1 timeout 300 python bench.py --no-cpu --no-e2e --no-gemm --no-extras --sustain-s 0 > $O/b1.json 2>$O/b1.err
python -c "
import json
d=json.loads(open('$O/b1.json').read().strip().splitlines()[-1])
print('L2_AHEAD=1: value', round(d['value'],1))" | tee -a $O/result.txt
timeout 120 python tools/xop_bench.py > $O/xop_bench.txt 2>&1; head -4 $O/xop_bench.txt
