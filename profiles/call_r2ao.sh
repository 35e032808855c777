#!/bin/bash
# Round 2, call ao (2 GPUs): the whole -m gpu suite as the driver runs it, on a box where the multi-GPU test does not skip
set -u
O=gpurun_out/r2ao
mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q --timeout 900 -p no:cacheprovider -rs > $O/pytest_gpu_2gpus.log 2>&1; echo "pytest rc=$? $(tail -1 $O/pytest_gpu_2gpus.log)"
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.txt 2>&1; echo "smoke rc=$?"; cat $O/smoke.txt | tail -4
