#!/bin/bash
# Round 2, call az: time line of the resident decode server's jobs (host-tensor calls)
set -u
O=gpurun_out/r2az
mkdir -p $O
LAMM_SERVER_TRACE=1 timeout 90 python tools/trace_server.py > $O/trace_server.txt 2>&1
LAMM_SERVER_TRACE=1 timeout 90 python tools/trace_server.py --m 4096 --k 11008 >> $O/trace_server.txt 2>&1
cat $O/trace_server.txt | tail -40
