#!/bin/bash
# Round 2, call bf: Q5_1 with A in TMEM reads stale raw weights on some stages -- which change removes it?
O=gpurun_out/r2bf; mkdir -p $O
for v in 4 5; do
  echo "== variant $v" | tee -a $O/dbg.txt
  LAMM_LIB=$PWD/build_variants/liblammcu_dbg$v.so timeout 200 python tools/dbg_atm.py 2>&1 | grep "^rep" | tee -a $O/dbg.txt
done
