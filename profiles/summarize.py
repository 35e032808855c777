#!/usr/bin/env python
"""Turn the raw captures of profiles/capture_r1.sh (gpurun_out/<round>/) into the small tracked
summaries under profiles/:  python profiles/summarize.py gpurun_out/r1 r1"""
import csv
import io
import json
import os
import subprocess
import sys
from collections import defaultdict

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum", "launch__grid_size", "launch__block_size",
        "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem", "launch__waves_per_multiprocessor", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "sm__inst_executed_pipe_tensor.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_uniform.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_membar_per_issue_active.ratio"]


def launch_list(path, out):
    rows = []
    with open(path) as f:
        text = f.read()
    start = text.index('"ID"')
    for r in csv.DictReader(io.StringIO(text[start:])):
        if r["Metric Name"] == "gpu__time_duration.sum":
            v = float(r["Metric Value"].replace(",", ""))
            if r["Metric Unit"].startswith("us"):
                v *= 1e3
            elif r["Metric Unit"].startswith("ms"):
                v *= 1e6
            rows.append((r["Kernel Name"], r["Grid Size"], r["Block Size"], v))
    agg = defaultdict(lambda: [0, 0.0, 1e30, 0.0])
    for name, grid, block, ns in rows:
        a = agg[(name, grid, block)]
        a[0] += 1; a[1] += ns; a[2] = min(a[2], ns); a[3] = max(a[3], ns)
    tot = sum(a[1] for a in agg.values())
    with open(out, "w") as f:
        f.write("# ncu launch list (--metrics gpu__time_duration.sum --clock-control none): per-launch times are\n"
                "# cold-cache and serialised; the kernel's SHARE of the step is what must agree with bench.py\n")
        f.write("kernel,grid,block,launches,total_us,avg_us,min_us,max_us,share\n")
        for (name, grid, block), a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write('"%s","%s","%s",%d,%.1f,%.3f,%.3f,%.3f,%.4f\n' % (name, grid, block, a[0], a[1] / 1e3, a[1] / a[0] / 1e3,
                                                                  a[2] / 1e3, a[3] / 1e3, a[1] / tot))
    return len(rows)


def rep_summary(rep, out):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw[raw.index('"ID"'):])))
    hdr, units, data = rows[0], rows[1], rows[2:]
    idx = {h: i for i, h in enumerate(hdr)}
    with open(out, "w") as f:
        f.write("# from %s (ncu --set full --clock-control none --import-source on), one column per captured launch\n" % os.path.basename(rep))
        f.write("metric,unit," + ",".join("launch%d" % i for i in range(len(data))) + "\n")
        for k in ["Kernel Name", "Grid Size", "Block Size"]:
            f.write("%s,," % k + ",".join('"%s"' % d[idx[k]] for d in data) + "\n")
        for k in KEYS:
            if k in idx:
                f.write("%s,%s," % (k, units[idx[k]]) + ",".join(d[idx[k]].replace(",", "") for d in data) + "\n")
    return len(data)


def main():
    src, tag = sys.argv[1], sys.argv[2]
    here = os.path.dirname(os.path.abspath(__file__))
    n = launch_list(os.path.join(src, "launches.csv"), os.path.join(here, "%s_launches_bench.csv" % tag))
    print("launch list:", n, "launches")
    for f in sorted(os.listdir(src)):
        if f.endswith(".ncu-rep"):
            k = rep_summary(os.path.join(src, f), os.path.join(here, "%s_ncu_%s.csv" % (tag, f[:-8])))
            print(f, k, "launches")
    for f in ("bench.json", "sweep_decode_fused.jsonl", "sweep_gemm_n512.jsonl", "sweep_gemm_13b_n2048.jsonl", "pytest_gpu.log"):
        p = os.path.join(src, f)
        if os.path.exists(p):
            with open(p) as a, open(os.path.join(here, "%s_%s" % (tag, f)), "w") as b:
                b.write(a.read())


if __name__ == "__main__":
    main()
