#!/bin/bash
# Round 2, call at (2 GPUs): in-kernel stamps of a chain of row-sharded decode launches
set -u
O=gpurun_out/r2at
mkdir -p $O
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29731 tools/trace_sharded.py > $O/trace_sharded_n2.txt 2>&1; echo rc=$?; grep -v "OMP_NUM\|\*\*\*\*" $O/trace_sharded_n2.txt | tail -16
