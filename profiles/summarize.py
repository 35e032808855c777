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

KEYS = ["gpu__time_duration.sum", "sm__cycles_elapsed.avg.per_second", "sm__cycles_elapsed.avg", "sm__cycles_active.avg",
        "dram__bytes_read.sum", "dram__bytes_write.sum",
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
    with open(path) as f:
        text = f.read()
    start = text.index('"ID"')
    per = {}                                       # launch id -> [name, grid, block, ns, dram read B, dram write B]
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    for r in csv.DictReader(io.StringIO(text[start:])):
        e = per.setdefault(r["ID"], [r["Kernel Name"], r["Grid Size"], r["Block Size"], 0.0, 0.0, 0.0])
        v = float(r["Metric Value"].replace(",", ""))
        u = r["Metric Unit"]
        if r["Metric Name"] == "gpu__time_duration.sum":
            e[3] = v * (1e3 if u.startswith("us") else 1e6 if u.startswith("ms") else 1.0)
        elif r["Metric Name"] == "dram__bytes_read.sum":
            e[4] = v * scale.get(u, 1.0)
        elif r["Metric Name"] == "dram__bytes_write.sum":
            e[5] = v * scale.get(u, 1.0)
    agg = defaultdict(lambda: [0, 0.0, 1e30, 0.0, 0.0, 0.0])
    for name, grid, block, ns, rd, wr in per.values():
        a = agg[(name, grid, block)]
        a[0] += 1; a[1] += ns; a[2] = min(a[2], ns); a[3] = max(a[3], ns); a[4] += rd; a[5] += wr
    tot = sum(a[1] for a in agg.values())
    with open(out, "w") as f:
        f.write("# ncu launch list (--metrics gpu__time_duration.sum,dram__bytes_{read,write}.sum --clock-control none):\n"
                "# per-launch times are cold-cache and serialised; the kernel's SHARE of the step is what must agree\n")
        f.write("kernel,grid,block,launches,total_us,avg_us,min_us,max_us,share,avg_dram_read_MB,avg_dram_write_MB\n")
        for (name, grid, block), a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write('"%s","%s","%s",%d,%.1f,%.3f,%.3f,%.3f,%.4f,%.3f,%.3f\n' % (
                name, grid, block, a[0], a[1] / 1e3, a[1] / a[0] / 1e3, a[2] / 1e3, a[3] / 1e3, a[1] / tot,
                a[4] / a[0] / 1e6, a[5] / a[0] / 1e6))
        n = sum(a[0] for a in agg.values())
        f.write("# all kernels: %d launches, avg dram read %.3f MB + write %.3f MB per launch\n" % (
            n, sum(a[4] for a in agg.values()) / n / 1e6, sum(a[5] for a in agg.values()) / n / 1e6))
    return len(per)


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
    for f in ("bench.json", "bench_nobatch.json", "bench_reference.json", "sweep_decode_fused.jsonl", "sweep_gemm_n512.jsonl",
              "sweep_gemm_13b_n2048.jsonl", "sweep_gemm_30b_q5_0_n512.jsonl", "pytest_gpu.log", "trace_decode.txt",
              "trace_server.txt", "hostlat.txt"):
        p = os.path.join(src, f)
        if os.path.exists(p):
            with open(p) as a, open(os.path.join(here, "%s_%s" % (tag, f)), "w") as b:
                b.write(a.read())


if __name__ == "__main__":
    main()
