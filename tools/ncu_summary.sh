#!/bin/bash
# usage: scratch_ncu.sh report.ncu-rep
ncu -i "$1" --page raw --csv 2>/dev/null | python -c "
import csv,sys
rows=list(csv.reader(sys.stdin))
hdr=rows[0]; units=rows[1]
pat=['gpu__time_duration.sum','dram__bytes_read.sum','dram__bytes_write.sum','dram__throughput.avg.pct','sm__warps_active.avg.pct','launch__registers','launch__grid_size','launch__block_size','launch__occupancy_limit','sm__throughput.avg.pct','bank_conflicts','smsp__cycles_active.avg','sm__cycles_elapsed.max','launch__shared_mem_per_block_dynamic','smsp__inst_executed.sum','issue_active','launch__waves','smsp__average_warp','smsp__warp_issue_stalled','smsp__average_warps_issue_stalled','l1tex__data_pipe_lsu_wavefronts_mem_shared','sm__pipe_tensor','achieved_occupancy','sm__inst_executed_pipe']
r=rows[2]
for i,h in enumerate(hdr):
    if any(p in h for p in pat):
        v=r[i]
        try:
            if float(v.replace(',',''))==0: continue
        except: pass
        print(h, v, units[i])
"
