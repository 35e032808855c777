#!/usr/bin/env python
"""Static evidence from the built library, no GPU needed:  python tools/static_report.py > profiles/r2_static_sass.md

Per kernel of la-llama.cpp_b200/liblammcu.so: registers / stack / shared memory (cuobjdump
--dump-resource-usage) and how often the SASS mnemonics that prove the Blackwell paths occur
(/opt/skills/guides/B200_PROFILING.md: tcgen05.mma -> UTC*MMA (kind::f16 UTCHMMA, kind::i8 UTCIMMA; cta_group::2 -> .2CTA), tcgen05.ld/st -> LDTM/STTM,
TMA -> UTMALDG / UTMASTG / UBLKCP; dp4a -> IDP.4A; mbarrier -> SYNCS; griddepcontrol -> ACQBULK /
PREEXIT family).  Template instantiations are grouped by kernel name with min..max over the group."""
import os
import re
import subprocess
import sys
from collections import OrderedDict, defaultdict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "la-llama.cpp_b200", "liblammcu.so")
MNEMONICS = ["UTCHMMA", "UTCIMMA", ".2CTA", "UTCQMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "UBLKPF", "IDP.4A", "IDP.2A",
             "SYNCS", "ACQBULK", "PREEXIT", "HMMA", "IMMA", "STL", "LDL"]


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
    return dict(zip(names, out))


def base_name(d):
    d = re.sub(r"^void ", "", d)
    return re.sub(r"<.*", "", d.split("(")[0])


def main():
    res = subprocess.run(["cuobjdump", "--dump-resource-usage", LIB], capture_output=True, text=True).stdout
    usage = {}
    cur = None
    for line in res.splitlines():
        m = re.search(r"Function (\S+):", line)
        if m:
            cur = m.group(1)
            continue
        if cur and "REG:" in line:
            usage[cur] = {k: int(v) for k, v in re.findall(r"(REG|STACK|SHARED|LOCAL):(\d+)", line)}
            cur = None
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    counts = defaultdict(lambda: defaultdict(int))
    ninstr = defaultdict(int)
    cur = None
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            continue
        m = re.search(r"/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if cur and m:
            op = m.group(1)
            ninstr[cur] += 1
            for k in MNEMONICS:
                if op == k or op.startswith(k + ".") or (k.startswith(".") and k in op):
                    counts[cur][k] += 1
    names = demangle(sorted(usage))
    groups = OrderedDict()
    for mangled in sorted(usage, key=lambda n: names[n]):
        groups.setdefault(base_name(names[mangled]), []).append(mangled)

    def rng(vals):
        lo, hi = min(vals), max(vals)
        return str(lo) if lo == hi else "%d..%d" % (lo, hi)
    print("# Static report of liblammcu.so (sm_100a) — `python tools/static_report.py`\n")
    print("Registers / stack bytes / static shared bytes per thread-block from `cuobjdump --dump-resource-usage`;")
    print("SASS mnemonic counts from `cuobjdump -sass` (min..max over a kernel's template instantiations).")
    print("Dynamic shared memory is set at launch and not shown here.\n")
    used = [k for k in MNEMONICS if any(counts[f][k] for f in usage)]
    print("| kernel | inst. | regs | stack | SASS instr | " + " | ".join(used) + " |")
    print("|---|---|---|---|---|" + "---|" * len(used))
    for g, fns in groups.items():
        row = [g, str(len(fns)), rng([usage[f]["REG"] for f in fns]), rng([usage[f].get("STACK", 0) for f in fns]),
               rng([ninstr[f] for f in fns])]
        row += [rng([counts[f][k] for f in fns]) for k in used]
        print("| " + " | ".join(row) + " |")
    print("\nReading: `UBLKCP` = 1-D TMA bulk copies (decode / GEMV weight rings), `UTMALDG` = tensor-map TMA loads (GEMM")
    print("operands), `UTCHMMA` = `tcgen05.mma.kind::f16`, `UTCIMMA` = `tcgen05.mma.kind::i8` (the Q2_K GEMM), `.2CTA` = instructions of the `cta_group::2` pair form (MMA, TMA load, commit multicast, TMEM allocation), `LDTM` = `tcgen05.ld` (TMEM → registers in the epilogues / drain),")
    print("`IDP.4A` / `IDP.2A` = dp4a / dp2a of the decode and GEMV dot products, `SYNCS` = mbarrier operations,")
    print("`STL`/`LDL` = local-memory (spill) traffic.")
    return 0


if __name__ == "__main__":
    sys.exit(main())
