#!/bin/bash
# Round 2, call bu: the suite and the two bench arms at the round's last commit
set -u
O=gpurun_out/r2
mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q --timeout 300 -p no:cacheprovider -rs > $O/pytest_gpu.log 2>&1; echo "pytest rc=$? $(tail -1 $O/pytest_gpu.log)"
timeout 600 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 5 --warmup 3 > $O/bench_reference.json 2>&1; echo "reference arm rc=$?"
