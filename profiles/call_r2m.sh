#!/bin/bash
# Round 2, call m: the driver's bench invocation with the new line (gemm block, sustained, chain check, honest e2e)
set -u
O=gpurun_out/r2m
mkdir -p $O
( time timeout 600 python bench.py > $O/bench.json 2> $O/bench.err ) 2> $O/bench.time; echo "bench rc=$? $(grep real $O/bench.time)"; tail -3 $O/bench.err
python - <<'P'
import json
d = json.loads(open("gpurun_out/r2m/bench.json").read().strip().splitlines()[-1])
print("value", d["value"], "frac", d["roofline"]["frac"], "sustained", d["sustained"], "chain_check", d["chain_check"])
print("e2e", {k: d["e2e"][k] for k in ("value", "steps", "h2d_bytes_per_step", "distinct_host_weight_buffers", "results_checked_against_device_path")})
print("phases", d["e2e"]["per_call_phases"])
for c in d["gemm"]["cases"]: print("gemm", c["fmt"], c["M"], c["K"], c["N"], c["us"], c["TOPS"], c["frac_int8_peak_measured"])
print("gemm cpu", d["gemm"].get("cpu_baseline"))
print("cpu", d["cpu_baseline"]["value"], d["cpu_baseline"]["cores"], "clocks", d["clocks"])
P
