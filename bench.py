#!/usr/bin/env python
"""bench.py -- the LAMM hot path on B200: LLaMA-7B Q4_1 decode tokens/s through the
quantized mul_mat chain, with the GEMV's HBM roofline and the reference's
own CPU path timed beside it.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--fmt q4_1]

One "step" = one decoded token = the weight matmuls of one LLaMA-7B forward pass at
N=1: 32 layers x {wq, wk, wv, wo (4096x4096), gate, up (11008x4096), down (4096x11008)}
+ the output head (32000x4096), each one `activation quantize -> quantized GEMV`, run
back to back on one stream (captured once into a CUDA graph).  Attention / norm / rope /
sampling are not on the path and are not run (stated in `config`).  All 225 weight
matrices are distinct device buffers (4.13 GB for Q4_1 >> 126 MB L2), so every step
streams its weights from HBM.

value : tokens/s with inputs resident in HBM (CUDA events on the launching stream).
e2e   : tokens/s through the drop-in operator `lamm_mul_mat` (ggml_tensor structs over
        HOST buffers; per matmul H2D of X, kernels, D2H of C inside the timed region).
roofline : the GEMV kernel alone (pre-quantized activations), algorithmic bytes / time.
cpu_baseline / --impl reference : the UNMODIFIED reference (oracle/_ref/libref.so,
        lamm AVX2 backend) through ggml_graph_compute on the host cores.

With --gpus N > 1 (torchrun) every weight is row-sharded M/N per rank and the output
slices are all-gathered over NVLink after every matmul (strong scaling).
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# ---- the workload (SURVEY 8d "tok/s definition") ---------------------------------------
MODELS = {"7b": (4096, 11008, 32000, 32), "13b": (5120, 13824, 32000, 40), "30b": (6656, 17920, 32000, 60)}
MODEL = "7b"
H, F, V, LAYERS = MODELS[MODEL]
LAYER_SHAPES = HEAD_SHAPE = None


def set_model(name):
    """LLaMA-7B is the metric's model; 13B / 30B shapes (BASELINE.json configs 4-5) for the scaling runs."""
    global MODEL, H, F, V, LAYERS, LAYER_SHAPES, HEAD_SHAPE
    MODEL = name
    H, F, V, LAYERS = MODELS[name]
    LAYER_SHAPES = [("wq", H, H), ("wk", H, H), ("wv", H, H), ("wo", H, H),
                    ("gate", F, H), ("up", F, H), ("down", H, F)]          # (name, M, K)
    HEAD_SHAPE = ("head", V, H)


set_model("7b")
TYPE_IDS = {"f32": 0, "f16": 1, "q4_0": 2, "q4_1": 3, "q5_0": 6, "q5_1": 7, "q8_0": 8, "q2_k": 10,
            "q3_k": 11, "q4_k": 12, "q5_k": 13, "q6_k": 14}          # 11-14: SURVEY 8f-1 (sweeps only)
BLCK = {0: 1, 1: 1, 2: 32, 3: 32, 6: 32, 7: 32, 8: 32, 10: 256, 11: 256, 12: 256, 13: 256, 14: 256}
TSIZE = {0: 4, 1: 2, 2: 18, 3: 20, 6: 22, 7: 24, 8: 34, 10: 84, 11: 110, 12: 144, 13: 176, 14: 210}
VEC_DOT = {0: 0, 1: 1, 2: 8, 3: 9, 6: 8, 7: 9, 8: 8, 10: 15, 11: 15, 12: 15, 13: 15, 14: 15}
ACT_SIZE = {0: (1, 4), 1: (1, 2), 8: (32, 34), 9: (32, 36), 15: (256, 292)}


# dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernel, averaged over the chain's
# launches, from the committed `ncu --set full` capture (profiles/); None until measured for a format
# q4_1: profiles/r1_launches_bench.csv -- 903 launches of decode16_kernel<q4_1>, 32.075 MB read + 0.002 MB written on
# average against 32.09 MB algorithmic (ratio 0.9995: every weight byte crosses DRAM once, C and X stay in L2)
TRAFFIC_PER_LAUNCH = {"q4_1": 32.077e6}
# the persistent chain kernel: one launch per token (filled in from profiles/ once captured)
TRAFFIC_PER_CHAIN_LAUNCH = {}


def chain_ops():
    ops = []
    for _ in range(LAYERS):
        ops.extend(LAYER_SHAPES)
    ops.append(HEAD_SHAPE)
    return ops


def w_row_bytes(t, k):
    return (k // BLCK[t]) * TSIZE[t]


def gemv_bytes(t, m, k, n=1):
    """Algorithmic bytes of one mul_mat: W blocks + f32 X + f32 C."""
    return m * w_row_bytes(t, k) + n * k * 4 + n * m * 4


def gemv_kernel_bytes(t, m, k, n=1):
    """Algorithmic bytes of the GEMV kernel alone: W blocks + quantized X + f32 C."""
    blck, size = ACT_SIZE[VEC_DOT[t]]
    return m * w_row_bytes(t, k) + n * (k // blck) * size + n * m * 4


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


# ---- clocks during the timed region ---------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        if os.environ.get("LAMM_BENCH_NO_SAMPLER"):          # experiment knob: is the sampler perturbing the timed region?
            return
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons, pw = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            if len(r) < 7:
                continue
            try:
                sm.append(float(r[0])); mx.append(float(r[1])); pw.append(float(r[2]))
            except ValueError:
                continue
            for nme, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nme)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


# ---- synthetic quantized tensors, generated on the device -----------------------------
# variance of the zero-mean integer a weight element takes in the unit-gain recipe below
_QVAR = {2: 18.6667, 3: 21.25, 6: 82.6667, 7: 85.25, 8: 5461.25, 10: 97.0}


def synth_weight(torch, t, m, k, gen, device):
    """Random ggml blocks (SURVEY 8d 'alternative fast path': uniform random quant bytes, finite fp16
    scales), with the scales chosen so that the matrix has ZERO-MEAN elements of variance 1/K: a chain
    of such matmuls (each one's output is the next one's activation row) keeps |x| ~ 1 for hundreds of
    steps instead of overflowing.  Q4_1/Q5_1: m = -7.5 d / -15.5 d; Q4_0/Q5_0: nibbles in 1..15 so
    that q-8 / q-16 is symmetric; Q2_K: scales byte = s | s<<4 (s in 1..15), dmin = 1.5 d.
    Returns uint8 [m, row_bytes]."""
    if t == 0:
        return (torch.randn((m, k), generator=gen, device=device, dtype=torch.float32) * (k ** -0.5)).view(torch.uint8)
    if t == 1:                                           # F16 (sweeps only)
        return (torch.randn((m, k), generator=gen, device=device, dtype=torch.float32) * (k ** -0.5)).to(torch.float16).view(torch.uint8)
    kb, bs = k // BLCK[t], TSIZE[t]
    raw = torch.randint(0, 256, (m, kb, bs), generator=gen, device=device, dtype=torch.uint8)
    halves = raw.view(torch.int16)                       # [m, kb, bs/2]
    if t in (11, 12, 13, 14):                            # Q3_K..Q6_K (sweeps only): random fields, finite fp16 d / dmin
        d16 = (torch.rand((m, kb), generator=gen, device=device) * 0.009 + 0.001).to(torch.float16).view(torch.int16)
        if t in (12, 13):
            halves[:, :, 0] = d16
            halves[:, :, 1] = d16
        else:
            halves[:, :, bs // 2 - 1] = d16
        return raw.view(m, kb * bs)
    base = (k * _QVAR[t] * 1.0208) ** -0.5
    d = (torch.rand((m, kb), generator=gen, device=device) * 0.5 + 0.75) * base
    d16 = d.to(torch.float16)

    def nib_1_15(lo, hi):                                # both nibbles of every qs byte uniform in 1..15
        shape = (m, kb, hi - lo)
        a = torch.randint(1, 16, shape, generator=gen, device=device, dtype=torch.uint8)
        b = torch.randint(1, 16, shape, generator=gen, device=device, dtype=torch.uint8)
        raw[:, :, lo:hi] = a | (b << 4)
    if t == 2:                                           # q4_0 {d, qs[16]}
        halves[:, :, 0] = d16.view(torch.int16)
        nib_1_15(2, 18)
    elif t == 3:                                         # q4_1 {d, m, qs[16]}
        halves[:, :, 0] = d16.view(torch.int16)
        halves[:, :, 1] = (d16.float() * -7.5).to(torch.float16).view(torch.int16)
    elif t == 6:                                         # q5_0 {d, qh[4], qs[16]}
        halves[:, :, 0] = d16.view(torch.int16)
        nib_1_15(6, 22)
    elif t == 7:                                         # q5_1 {d, m, qh[4], qs[16]}
        halves[:, :, 0] = d16.view(torch.int16)
        halves[:, :, 1] = (d16.float() * -15.5).to(torch.float16).view(torch.int16)
    elif t == 8:                                         # q8_0 {d, qs[32]}
        halves[:, :, 0] = d16.view(torch.int16)
    elif t == 10:                                        # q2_K {scales[16], qs[64], d, dmin}
        sc = raw[:, :, 0:16]
        lo = sc & 0x0F
        lo = torch.where(lo == 0, torch.ones_like(lo), lo)
        sc.copy_(lo | (lo << 4))
        halves[:, :, 40] = d16.view(torch.int16)
        halves[:, :, 41] = (d16.float() * 1.5).to(torch.float16).view(torch.int16)
    return raw.view(m, kb * bs)


def host_threads():
    try:
        import psutil
        n = psutil.cpu_count(logical=False) or os.cpu_count() or 1
    except Exception:
        n = os.cpu_count() or 1
    try:
        n = min(n, len(os.sched_getaffinity(0)))
    except AttributeError:
        pass
    p = 1
    while p * 2 <= min(n, 64):   # power of two: must divide M (reference drops M % nth rows)
        p *= 2
    return p


# ---- the reference's own CPU path ------------------------------------------------------
def cpu_reference_run(t, iters, warmup, nth=None):
    """Times the unmodified reference (oracle/_ref/libref.so; falls back to the oracle port)
    on the host cores.  Per step: ONE timed ggml_graph_compute of each distinct matmul shape of
    the token chain (plus the reference benchmark's untimed cache-thrash graph in between),
    extrapolated by the shapes' multiplicities: t_token = 32*(4 a + 2 b + c) + head."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from oracle_lib import Oracle, Ref
    nth = nth or host_threads()
    rng = np.random.default_rng(5)
    shapes = [("attn", H, H, 4 * LAYERS), ("gate_up", F, H, 2 * LAYERS), ("down", H, F, LAYERS), ("head", V, H, 1)]
    per = {}
    kind = "reference" if Ref.available() else "port"
    if kind == "reference":
        r = Ref()
        for name, m, k, _ in shapes:
            x = rng.standard_normal((1, k)).astype(np.float32)
            w = rng.integers(0, 256, (m, w_row_bytes(t, k)), dtype=np.uint8)
            fix_scales_host(w, t, k)
            per[name] = np.array(r.bench(t, m, 1, k, nth, warmup + iters, wq=w, x=x))[warmup:] * 1e-6
    else:
        o = Oracle()
        for name, m, k, _ in shapes:
            x = rng.standard_normal((1, k)).astype(np.float32)
            w = rng.integers(0, 256, (m, w_row_bytes(t, k)), dtype=np.uint8)
            fix_scales_host(w, t, k)
            ts = []
            for _ in range(warmup + iters):
                t0 = time.perf_counter()
                o.mul_mat_mt(t, w, x, nth)
                ts.append(time.perf_counter() - t0)
            per[name] = np.array(ts[warmup:])
    tok = sum(per[name] * mult for name, _, _, mult in shapes)        # seconds per token, per step
    return {"sec_per_token": tok, "kind": kind, "cores": nth,
            "sample": "per step one timed ggml_graph_compute (INIT quantize + lamm_mul_mat, N=1) of each "
                      "distinct chain shape (HxH, FxH, HxF, VxH of the model), extrapolated "
                      "x(4L, 2L, L, 1); cache-thrash graph between timings as in la-benchmark-matmult",
            "per_shape_us": {n_: float(np.mean(v) * 1e6) for n_, v in per.items()}}


def fix_scales_host(w, t, k):
    """Give random host blocks sane fp16 scales (same recipe as synth_weight)."""
    if t == 0:
        w.view(np.float32)[:] = 0.02
        return
    kb, bs = k // BLCK[t], TSIZE[t]
    h = w.reshape(w.shape[0], kb, bs).view(np.uint16)
    rng = np.random.default_rng(1)
    sc = lambda: (rng.random((w.shape[0], kb), np.float32) * 0.029 + 0.001).astype(np.float16).view(np.uint16)
    if t in (2, 6, 8):
        h[:, :, 0] = sc()
    elif t in (3, 7):
        h[:, :, 0] = sc(); h[:, :, 1] = sc()
    elif t == 10:
        h[:, :, 40] = sc(); h[:, :, 41] = sc()


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    t = TYPE_IDS[args.fmt]
    r = cpu_reference_run(t, args.steps, args.warmup)
    sec = float(np.mean(r["sec_per_token"]))
    val = 1.0 / sec
    line = {
        "impl": "reference", "metric": "llama%s_%s_decode_matmul_chain_tokens_per_s" % (MODEL, args.fmt),
        "value": val, "unit": "tok/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "i8 x u4 -> i32, f32 scale-accumulate (AVX2 maddubs)", "data": "synthetic",
        "config": workload_config(args.fmt, 1),
        "cpu_baseline": {"value": val, "unit": "tok/s", "cores": r["cores"], "kind": r["kind"],
                         "sample": r["sample"], "per_shape_us": r["per_shape_us"]},
        "e2e": {"value": val, "unit": "tok/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


def workload_config(fmt, world, collective="fused", ll=False):
    ops = chain_ops()
    cfg = {"workload": "LLaMA-%s %s decode (N=1) weight-matmul chain: %d x {4x %dx%d, 2x %dx%d, 1x %dx%d} + head %dx%d; "
                        "fused activation-quantize + quantized GEMV per matmul, data dependent (each output row is a "
                        "later matmul's activation row); attention/norm/rope/sampling excluded (not on the path)"
                        % (MODEL.upper(), fmt.upper(), LAYERS, H, H, F, H, H, F, V, H),
            "format": fmt, "n_tokens_per_step": 1, "matmuls_per_step": len(ops),
            "l2_defeat": "%d distinct weight buffers per step (%.2f GB) >> 126 MB L2" % (
                len(ops), sum(m * w_row_bytes(TYPE_IDS[fmt], k) for _, m, k in ops) / 1e9),
            "parallelism": "single GPU",
            "activation_vectors": "LL {f32 value, u32 tag} pairs, consumer polls tags (no kernel-boundary wait)" if ll
                                  else "plain f32 (ggml layout)"}
    if world > 1:
        cfg["parallelism"] = ("row-sharded M/%d per rank; C replicated on every rank by %s" % (
            world, "the decode kernel's NVLink peer-store epilogue ({value, tag} 64-bit stores, consumer polls; "
                   "no collective call)" if collective == "fused" else "ncclAllGather after every matmul"))
    return cfg


# ---- the CUDA path ---------------------------------------------------------------------
def run_cuda(args):
    import torch
    import lamm_b200 as L
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("--gpus %d needs torchrun (one rank per GPU)" % args.gpus)
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (there is no CPU fallback)"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    L.init(local)
    lib = L.lib()
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)

    t = TYPE_IDS[args.fmt]
    vt = VEC_DOT[t]
    ops = chain_ops()
    gen = torch.Generator(device=dev)
    gen.manual_seed(1234)

    def shard(m):                      # contiguous rows [lo, hi) of this rank
        per = (m + world - 1) // world
        lo = min(m, rank * per)
        return lo, min(m, lo + per)

    # weights: the rank's row shard of every matrix, distinct buffers
    weights = []
    for name, m, k in ops:
        lo, hi = shard(m)
        weights.append(synth_weight(torch, t, hi - lo, k, gen, dev))
    x0 = torch.randn((1, H), generator=gen, device=dev)          # the token's input row (layer 0)
    act_blck, act_size = ACT_SIZE[vt]
    xq_ws = torch.zeros(((F // act_blck) * act_size + 64,), dtype=torch.uint8, device=dev)
    fused = ((world > 1 and args.collective == "fused") or args.chain == "ll") and t != 0
    for m in (H, F, V):
        assert m % world == 0, "row shards must be equal for the in-place all-gather"

    # activation / output vectors of the chain.  The chain is DATA DEPENDENT, as in the model:
    #   q,k,v = W{q,k,v} x;  o = Wo v;  g = Wgate o;  u = Wup o;  x' = Wdown u;  logits = Whead x_last
    # (attention / SiLU.mul / norm are not on the path, so v and u stand in for their results).
    comm = None
    if fused:
        # peer-store epilogue (SURVEY 8e "preferred"): every rank's decode kernel writes its rows of C as
        # {value, tag} pairs straight into the full vector of every rank over NVLink; the consumer launch
        # polls the tags.  No collective call, fence or flag on the data path.
        from lamm_b200 import sharding
        comm = sharding.Comm(L, dist)
        mk = comm.vector
    else:
        mk = lambda m: torch.zeros((1, m), device=dev)
    buf = {"x0": mk(H), "x1": mk(H), "q": mk(H), "k": mk(H), "v": mk(H), "o": mk(H), "g": mk(F), "u": mk(F),
           "logits": mk(V)}
    plan = []                                   # (weight index, M, K, input key or None for x0 tensor, output key)
    wi = 0
    for layer in range(LAYERS):
        xin_key = None if layer == 0 else "x%d" % (layer % 2)
        xout_key = "x%d" % ((layer + 1) % 2)
        for name, src, dst in (("wq", xin_key, "q"), ("wk", xin_key, "k"), ("wv", xin_key, "v"), ("wo", "v", "o"),
                               ("gate", "o", "g"), ("up", "o", "u"), ("down", "u", xout_key)):
            _, m, k = ops[wi]
            plan.append((wi, m, k, src, dst))
            wi += 1
    plan.append((wi, V, H, "x%d" % (LAYERS % 2), "logits"))

    # sibling matmuls that share an activation row (q/k/v, gate/up) go out as ONE launch
    # (lammcu_mul_mat_batch_dev) unless --no-batch; same 225 matmuls, fewer dependency waits
    groups = []
    for item in plan:
        if (groups and args.batch and (fused or world == 1) and t != 0 and groups[-1][0][3] == item[3]
                and groups[-1][0][1:3] == item[1:3] and len(groups[-1]) < 4):
            groups[-1].append(item)
        else:
            groups.append([item])

    chain = None
    if args.chain == "persistent" and world == 1 and t != 0:
        # ONE persistent launch per token (lammcu_chain_*): the same groups, separated by an in-kernel grid barrier
        chain = L.Chain(t, [([weights[g[0]] for g in grp], x0 if grp[0][3] is None else buf[grp[0][3]],
                             [buf[g[4]] for g in grp]) for grp in groups])

    def one_token():
        if chain is not None:
            chain.run()
            return
        if args.batch and fused:
            for grp in groups:
                xin = x0 if grp[0][3] is None else buf[grp[0][3]]
                if len(grp) == 1:
                    wi, m, k, src, dst = grp[0]
                    comm.mul_mat(t, weights[wi], xin, buf[dst], m, k)
                else:
                    comm.mul_mat_batch(t, [weights[g[0]] for g in grp], xin, [buf[g[4]] for g in grp],
                                       [g[1] for g in grp], grp[0][2])
            return
        if args.batch and not fused and world == 1 and t != 0:
            for grp in groups:
                xin = x0 if grp[0][3] is None else buf[grp[0][3]]
                if len(grp) == 1:
                    wi, m, k, src, dst = grp[0]
                    L.mul_mat(t, weights[wi], xin, out=buf[dst])
                else:
                    L.mul_mat_batch(t, [weights[g[0]] for g in grp], xin, [buf[g[4]] for g in grp])
            return
        for wi, m, k, src, dst in plan:
            w = weights[wi]
            rows = w.shape[0]
            xin = x0 if src is None else buf[src]
            if fused:
                comm.mul_mat(t, w, xin, buf[dst], m, k)
                continue
            wv = w if t != 0 else w.view(torch.float32)
            lo = rank * rows
            L.mul_mat(t, wv, xin, out=buf[dst][:, lo:lo + rows])
            if world > 1:
                dist.all_gather_into_tensor(buf[dst].view(-1), buf[dst][0, lo:lo + rows])   # in place

    def capture(fn):
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            for _ in range(2):
                fn()
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        lib.lammcu_launch_count_reset()
        with torch.cuda.graph(g):
            fn()
        return g, int(lib.lammcu_launch_count())

    def timed(g, steps, warmup):
        for _ in range(warmup):
            g.replay()
        torch.cuda.synchronize()
        if dist:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            g.replay()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        if dist:
            tt = torch.tensor([ms], device=dev)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            dist.barrier()
            ms = float(tt.item())
        return ms

    g_full, launches_per_token = capture(one_token)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()               # runs across warm-up, the timed region, the sustained run, the gemm and e2e legs
    ms = timed(g_full, args.steps, max(args.warmup, 3))
    tok_s = args.steps / (ms * 1e-3)
    # the same graph for >= 2 s back to back: what the part sustains once clocks and power have settled
    sustained = None
    if args.sustain_s > 0:
        n_sus = max(args.steps, int(args.sustain_s * tok_s))
        ms_sus = timed(g_full, n_sus, 0)
        sustained = {"value": n_sus / (ms_sus * 1e-3), "unit": "tok/s", "steps": n_sus, "seconds": ms_sus * 1e-3}
    # the chain's final output must be finite (unit-gain weights) and, with N>1, no poll may have timed out
    logits = comm.unpack(torch, buf["logits"]) if fused else buf["logits"].view(-1)
    torch.cuda.synchronize()
    if not bool(torch.isfinite(logits).all().item()):
        raise SystemExit("bench: non-finite logits at the end of the chain")
    if fused and comm.timed_out():
        raise SystemExit("bench: a peer poll timed out")
    logit_rms = float(logits.float().pow(2).mean().sqrt().item())
    # ... and RIGHT: the chain once more without CUDA graph, sibling batching, PDL overlap or LL vectors -- one plain
    # lammcu_mul_mat_dev per matmul on fresh buffers -- must give the same logits (a stale tag, a WAR on the reused
    # activation vectors under graph replay or a mis-addressed sibling batch would still be finite)
    chain_check = None
    if world == 1:
        ref_buf = {k_: torch.zeros((1, {"g": F, "u": F, "logits": V}.get(k_, H)), device=dev) for k_ in buf}
        for wi, m, k, src, dst in plan:
            wv = weights[wi] if t != 0 else weights[wi].view(torch.float32)
            L.mul_mat(t, wv, x0 if src is None else ref_buf[src], out=ref_buf[dst])
            torch.cuda.synchronize()
        want = ref_buf["logits"].view(-1)
        err = float((logits.float() - want).abs().max().item())
        tol = 1e-5 * float(np.sqrt(H)) * float(want.abs().max().item()) * 8     # 225 chained matmuls: a few ulps of drift
        chain_check = {"max_abs_err": err, "tol": tol, "vs": "un-batched, un-graphed lammcu_mul_mat_dev per matmul"}
        if not err <= tol:
            raise SystemExit("bench: the timed chain's logits differ from the plain chain: %g > %g" % (err, tol))

    # roofline of the dominant kernel (decode16_kernel, the ONLY kernel of the step at N=1): the
    # algorithmic bytes of this rank's launches (W blocks + f32 X + f32 C) over the average
    # launch duration measured above with CUDA events
    peak, peak_src = peaks()
    kern_bytes = sum(gemv_bytes(t, shard(m)[1] - shard(m)[0], k) for _, m, k in ops)
    achieved = kern_bytes * args.steps / (ms * 1e-3) / 1e9
    path_bytes = sum(gemv_bytes(t, m, k) for _, m, k in ops)
    launches_q = launches_per_token

    # e2e leg: the drop-in operator on HOST buffers (rank 0 / N=1 only does the H2D/D2H variant;
    # with N>1 every rank runs it on its shard)
    if args.host_mode is not None:
        L.lib().lammcu_set_host_mode(args.host_mode)
    e2e = None
    if not args.no_e2e and world == 1:
        e2e = e2e_run(torch, L, t, ops, weights, shard, max(50, args.steps // 4), dist, dev, rank, world)
    elif not args.no_e2e:
        # ggml needs the whole C on the caller's host after every matmul; the row-sharded operator path returns each
        # rank's slice only, so there is no honest end-to-end number at N > 1: it is left out rather than reported
        # on 1/N of the result
        e2e = {"value": None, "unit": "tok/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0,
               "unavailable": "the host operator is single-GPU; the sharded path is device-resident (value)"}

    gemm = None
    if world == 1 and rank == 0 and not args.no_gemm:
        gemm = gemm_block(torch, L, dev, args)
    extras = None
    if not args.no_extras and t != 0:
        if world > 1 and comm is None:
            from lamm_b200 import sharding
            comm = sharding.Comm(L, dist)
        extras = scale_extras(torch, L, dist, dev, rank, world, comm if world > 1 else None, args)
    clocks = sampler.stop() if rank == 0 else None
    if rank != 0:
        return finish(dist)

    cpu = None
    if world == 1 and not args.no_cpu:
        try:
            r = cpu_reference_run(t, 8, 2)
            # median over the 8 samples: this leg shares the host with whatever the legs before it left running (one
            # capture had the down-projection shape at 5x its usual time on average and 11.5 tok/s in the reference arm)
            cpu = {"value": 1.0 / float(np.median(r["sec_per_token"])), "unit": "tok/s", "cores": r["cores"],
                   "kind": r["kind"], "sample": r["sample"] + "; median of 8 samples", "per_shape_us": r["per_shape_us"]}
        except Exception as e:      # the checker library is absent: say so, do not fake a number
            cpu = {"value": None, "unit": "tok/s", "cores": 0, "kind": "unavailable", "sample": str(e)}

    line = {
        "metric": "llama%s_%s_decode_matmul_chain_tokens_per_s" % (MODEL, args.fmt),
        "value": tok_s, "unit": "tok/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "u4/u5/i8 weights x i8 activations -> i32 (dp4a), f32 scale-accumulate",
        "data": "synthetic", "config": dict(workload_config(args.fmt, world, "fused" if fused else "nccl", fused),
                                             launches_per_step=launches_per_token,
                                             sibling_batching=bool(args.batch and (fused or world == 1) and t != 0)),
        "path_gbs": path_bytes * args.steps / (ms * 1e-3) / 1e9,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": (TRAFFIC_PER_CHAIN_LAUNCH.get(args.fmt) if chain is not None else TRAFFIC_PER_LAUNCH.get(args.fmt)) if world == 1 else None,
                     "kernel": "lamm::decode_chain_kernel<%s>" % args.fmt if chain is not None else "lamm::decode16_kernel<%s>" % args.fmt,
                     "peak_source": peak_src,
                     "how": "the timed chain itself (every launch of the step is this kernel, activation quantize "
                            "fused), CUDA events; bytes = W blocks + f32 X + f32 C per launch; %d launches/step, "
                            "avg %.2f us and %.2f MB per launch"
                            % (launches_q, ms * 1e3 / args.steps / max(launches_q, 1),
                               kern_bytes / max(launches_q, 1) / 1e6)},
        "sustained": sustained, "chain_check": chain_check,
        "gemm": gemm,
        "multi_gpu_workloads": extras,
        "cpu_baseline": cpu,
        "e2e": e2e,
        "gpu_launches": launches_per_token * args.steps, "logits_rms": logit_rms,
        "clocks": clocks,
    }
    print(json.dumps(line))
    return finish(dist)


# measured on this pool's B200s with tools/i8peak.cu (tcgen05.mma kind::i8 / kind::f16, M128 x N256, operands resident
# in shared memory, all 148 SMs; profiles/r2g/i8peak.txt, re-measured in profiles/r2k/): the GEMM fractions' denominators
INT8_TOPS_MEASURED = 4563.3
F16_TOPS_MEASURED = 2220.9


def gemm_block(torch, L, dev, args):
    """The prefill third of BASELINE.json's metric ("GEMM TOPS vs roofline") in the driver's record: the two prefill
    configurations BASELINE names that fit one GPU -- LLaMA-13B Q4_1 N = 2048 (13824 x 5120, config 4) and LLaMA-7B
    Q2_K N = 512 (11008 x 4096, config 3) -- through lammcu_mul_mat_dev (activation quantize + operand prep + tcgen05
    GEMM, all inside the timed region), a ring of distinct weight matrices > L2 replayed from one CUDA graph, CUDA
    events.  TOPS = 2 M N K / time, as the reference counts flops (src/la-benchmark-matmult.cpp:331-332).  Beside
    them the reference's own CPU path at config 1's GEMM points: its benchmark's fixed shape K = 11008 M = 4096
    N = 128 (src/la-benchmark-matmult.cpp:180-182) and 4096 x 4096 x 512."""
    gen = torch.Generator(device=dev)
    gen.manual_seed(99)
    out = {"unit": "TOPS", "int8_peak_tops_measured": INT8_TOPS_MEASURED, "f16_peak_tflops_measured": F16_TOPS_MEASURED,
           "peak_source": "tools/i8peak.cu on this pool (profiles/r2g/i8peak.txt): an MMA stream on constant operands "
                          "resident in shared memory, which never meets the board's power cap", "cases": []}
    # the driver's dense figure: cuBLAS bf16 8192^3 on random data, burst / held for seconds under the 1 kW cap.  A
    # tensor-bound kernel on real data is power-bound well below the constant-operand stream above (profiles/r2bd: the
    # pair GEMM runs at 1485-1580 MHz of 1965 with sw_power_cap active), so the ISSUED fp16 tensor rate (3 MMAs per
    # product for the 32-element formats) is also reported against this file's burst number
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            pk = json.load(f)
        bf16_burst, bf16_sust = float(pk["bf16_tflops"]), float(pk.get("bf16_tflops_sustained", 0.0)) or None
    except (OSError, KeyError, ValueError):
        bf16_burst, bf16_sust = 1590.0, None                     # B200_PROFILING.md fallback
    out["bf16_dense_tflops_file"] = {"burst": bf16_burst, "sustained": bf16_sust,
                                     "source": "MEASURED_PEAKS.json (cuBLAS bf16)" if bf16_sust else "fallback (B200_PROFILING.md)"}
    for fmt, m, k, n, mmas, kind in (("q4_1", 13824, 5120, 2048, 3, "tcgen05 kind::f16 cta_group::2 (CTA pairs, 256 x 256 tiles), 3 MMAs per product (scaled fp16 hi/lo operands)"),
                                     ("q2_k", 11008, 4096, 512, 0, "tcgen05 kind::i8 (scale-folded u8 x s8, exact integer sums)")):
        t = TYPE_IDS[fmt]
        wb = m * w_row_bytes(t, k)
        ring = max(2, min(16, (256 * (1 << 20) + wb - 1) // wb))
        ws = [synth_weight(torch, t, m, k, gen, dev) for _ in range(ring)]
        x = torch.randn((n, k), generator=gen, device=dev)
        c = torch.zeros((n, m), device=dev)

        def body():
            for w in ws:
                L.mul_mat(t, w, x, out=c)
        s_ = torch.cuda.Stream()
        s_.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s_):
            body()
        torch.cuda.current_stream().wait_stream(s_)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        L.lib().lammcu_launch_count_reset()
        with torch.cuda.graph(g):
            body()
        launches = int(L.lib().lammcu_launch_count()) // ring
        for _ in range(3):
            g.replay()
        torch.cuda.synchronize()
        reps = max(3, 200 // ring)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            g.replay()
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / (reps * ring)
        tops = 2.0 * m * n * k / us / 1e6
        if not bool(torch.isfinite(c).all().item()):
            raise SystemExit("bench gemm: non-finite C (%s)" % fmt)
        out["cases"].append({"fmt": fmt, "M": m, "K": k, "N": n, "us": round(us, 2), "TOPS": round(tops, 1),
                             "frac_int8_peak_measured": round(tops / INT8_TOPS_MEASURED, 4),
                             "frac_f16_peak_measured": round(tops / F16_TOPS_MEASURED, 4),
                             "tensor_path": kind, "launches_per_matmul": launches, "distinct_weights": ring,
                             "includes": "activation quantize + operand prep + GEMM"})
        if mmas:
            out["cases"][-1]["issued_f16_tflops"] = round(tops * mmas, 1)
            out["cases"][-1]["issued_frac_of_file_bf16_burst"] = round(tops * mmas / bf16_burst, 4)
        del ws, g
        torch.cuda.empty_cache()
    if not args.no_cpu:
        try:
            sys.path.insert(0, os.path.join(ROOT, "tests"))
            from oracle_lib import Ref
            if Ref.available():
                r = Ref()
                nth = host_threads()
                cpu = []
                for (m, k, n, why) in ((4096, 11008, 128, "la-benchmark-matmult's fixed shape"), (4096, 4096, 512, "config 1")):
                    us_ = float(np.mean(r.bench(TYPE_IDS["q4_1"], m, n, k, nth, 4)[1:]))
                    cpu.append({"fmt": "q4_1", "M": m, "K": k, "N": n, "us": round(us_, 1),
                                "gflops": round(2.0 * m * n * k / us_ / 1e3, 1), "cores": nth, "kind": "reference", "why": why})
                out["cpu_baseline"] = cpu
        except Exception as e:
            out["cpu_baseline"] = [{"kind": "unavailable", "why": str(e)}]
    return out


def scale_extras(torch, L, dist, dev, rank, world, comm, args):
    """The two multi-GPU workloads BASELINE.json names besides the 7B chain (configs 4 and 5), as extra keys of the line at
    every N so the driver's 1/2/4/8 run carries them:
      * LLaMA-30B Q5_0 decode chain (K = 6656, M = 17920 ...), weights row-sharded, fused LL epilogue: tok/s
      * LLaMA-13B Q4_1 N = 2048 prefill on 13824 x 5120, rows sharded: aggregate TOPS compute-only (each rank's slab
        into its own buffer) and compute + replicate (the peer-store epilogue leaves the full C on every rank)
    Device time, CUDA events, barrier on both sides, max over ranks."""
    out = {}

    def timed(fn, reps):
        for _ in range(2):
            fn()
        torch.cuda.synchronize()
        if dist:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        if dist:
            tt = torch.tensor([ms], device=dev)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            dist.barrier()
            ms = float(tt.item())
        return ms / reps

    def shard(m):
        per = (m + world - 1) // world
        lo = min(m, rank * per)
        return lo, min(m, lo + per)
    gen = torch.Generator(device=dev)
    gen.manual_seed(4321)
    # ---- 30B Q5_0 decode chain ----
    Hh, Ff, Vv, Ll = MODELS["30b"]
    t = TYPE_IDS["q5_0"]
    if all(m % world == 0 for m in (Hh, Ff, Vv)):
        def W(m, k):
            lo, hi = shard(m)
            return synth_weight(torch, t, hi - lo, k, gen, dev)
        layers = [dict(qkv=[W(Hh, Hh) for _ in range(3)], wo=W(Hh, Hh), gu=[W(Ff, Hh) for _ in range(2)], down=W(Hh, Ff))
                  for _ in range(Ll)]
        head = W(Vv, Hh)
        mk = comm.vector if comm is not None else (lambda m: torch.zeros((1, m), device=dev))
        b = {k_: mk(m) for k_, m in (("x0", Hh), ("x1", Hh), ("q", Hh), ("k", Hh), ("v", Hh), ("o", Hh), ("g", Ff), ("u", Ff), ("logits", Vv))}
        x_in = torch.randn((1, Hh), generator=gen, device=dev)

        def mm(ws, xin, outs, ms, k):
            if comm is not None:
                if len(ws) == 1:
                    comm.mul_mat(t, ws[0], xin, outs[0], ms[0], k)
                else:
                    comm.mul_mat_batch(t, ws, xin, outs, ms, k)
            elif len(ws) == 1:
                L.mul_mat(t, ws[0], xin, out=outs[0])
            else:
                L.mul_mat_batch(t, ws, xin, outs)

        def token():
            for li, ly in enumerate(layers):
                xin = x_in if li == 0 else b["x%d" % (li % 2)]
                mm(ly["qkv"], xin, [b["q"], b["k"], b["v"]], [Hh] * 3, Hh)
                mm([ly["wo"]], b["v"], [b["o"]], [Hh], Hh)
                mm(ly["gu"], b["o"], [b["g"], b["u"]], [Ff] * 2, Hh)
                mm([ly["down"]], b["u"], [b["x%d" % ((li + 1) % 2)]], [Hh], Ff)
            mm([head], b["x%d" % (Ll % 2)], [b["logits"]], [Vv], Hh)
        s_ = torch.cuda.Stream()
        s_.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s_):
            token()
            token()
        torch.cuda.current_stream().wait_stream(s_)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            token()
        ms = timed(g.replay, 20)
        logits = comm.unpack(torch, b["logits"]) if comm is not None else b["logits"].view(-1)
        torch.cuda.synchronize()
        ok = bool(torch.isfinite(logits).all().item()) and not (comm is not None and comm.timed_out())
        wbytes = sum(w_row_bytes(t, k) * m for m, k in [(Hh, Hh)] * 4 + [(Ff, Hh)] * 2 + [(Hh, Ff)]) * Ll + w_row_bytes(t, Hh) * Vv
        out["llama30b_q5_0_decode_chain"] = {"value": 1e3 / ms if ok else None, "unit": "tok/s", "ms_per_token": ms,
                                              "weight_gb_total": wbytes / 1e9, "aggregate_gbs": wbytes / ms / 1e6,
                                              "sharding": "rows / %d, LL peer-store epilogue" % world if world > 1 else "one GPU"}
        del layers, head, g
        torch.cuda.empty_cache()
    # ---- 13B Q4_1 N = 2048 prefill, rows sharded ----
    t = TYPE_IDS["q4_1"]
    m, k, n = 13824, 5120, 2048
    if m % world == 0:
        lo, hi = shard(m)
        w = synth_weight(torch, t, hi - lo, k, gen, dev)
        x = torch.randn((n, k), generator=gen, device=dev)
        c_local = torch.zeros((n, hi - lo), device=dev)
        ms_c = timed(lambda: L.mul_mat(t, w, x, out=c_local), 10)
        rec = {"M": m, "K": k, "N": n, "rows_per_gpu": hi - lo, "compute_only_tops": 2.0 * m * n * k / ms_c / 1e9, "compute_only_us": ms_c * 1e3}
        if comm is not None:
            cbuf = comm.result_buffer(n, m)
            ms_r = timed(lambda: comm.mul_mat_gemm(t, w, x, cbuf, m, n, k), 10)
            rec.update(compute_and_replicate_tops=2.0 * m * n * k / ms_r / 1e9, compute_and_replicate_us=ms_r * 1e3,
                       replicated_mb_per_gpu=n * m * 4 / 1e6)
        out["llama13b_q4_1_n2048_prefill_row_sharded"] = rec
    return out


def finish(dist):
    """Leave without tearing NCCL down: captured graphs still reference the communicator and
    destroy_process_group() can block on them; every rank has already passed the last barrier."""
    sys.stdout.flush()
    sys.stderr.flush()
    if dist:
        dist.barrier()
        os._exit(0)
    return 0


def e2e_run(torch, L, t, ops, weights, shard, steps, dist, dev, rank, world):
    """tokens/s through lamm_can_mul_mat + lamm_mul_mat with ggml_tensor structs over HOST memory:
    the call ggml.c:10858-10863 makes.  Weights are registered (uploaded) on first use, outside the
    timed region, exactly once -- they are immutable model state; X goes H2D and C comes D2H inside
    every call."""
    lib = L.lib()
    # host copies of ALL the chain's weight matrices (225 distinct host buffers, 4.13 GB for 7B Q4_1): the operator's
    # registry is keyed by the host pointer, so the device then really holds 225 distinct buffers and every call
    # streams its weights from HBM, as config.l2_defeat says
    idx_for = list(range(len(ops)))
    host_w = {i: weights[i].cpu().numpy() for i in idx_for}
    # activation rows as a ggml graph has them: one src1 tensor per producer -- wq/wk/wv share the
    # attention-norm output, gate/up share the ffn-norm output, wo / down / head have their own
    rng_x = np.random.default_rng(3)
    x_of = {"wq": "attn_in", "wk": "attn_in", "wv": "attn_in", "wo": "attn_out", "gate": "ffn_in", "up": "ffn_in",
            "down": "ffn_act", "head": "final"}
    x_dim = {"attn_in": H, "attn_out": H, "ffn_in": H, "ffn_act": F, "final": H}
    xh = {n_: rng_x.standard_normal((1, k_)).astype(np.float32) for n_, k_ in x_dim.items()}
    ch = {m: np.zeros((1, m), np.float32) for m in (H, F, V)}
    calls = []
    keep = []
    for (name, m, k), wi in zip(ops, idx_for):
        w = host_w[wi]
        rows = w.shape[0]
        src0 = L.make_tensor_2d(t, k, rows, w.ctypes.data, w.strides[0])
        xr = xh[x_of[name]]
        src1 = L.make_tensor_2d(L.F32, k, 1, xr.ctypes.data, xr.strides[0])
        dst = L.make_tensor_2d(L.F32, rows, 1, ch[m].ctypes.data)
        dst.src[0] = C.pointer(src0)
        dst.src[1] = C.pointer(src1)
        p = L.GgmlComputeParams(L.TASK_COMPUTE, 0, 1, 0, None)
        keep.append((src0, src1, dst, p))
        calls.append((C.byref(p), C.byref(dst)))
    can, mm = lib.lamm_can_mul_mat, lib.lamm_mul_mat
    # the same calls from a C loop (tools/bench_driver.c), as ggml makes them; Python loop if it was not built
    drv = None
    drv_path = os.path.join(ROOT, "tools", "libbench_driver.so")
    if os.path.exists(drv_path) and not os.environ.get("LAMM_BENCH_PYLOOP"):
        drv = C.CDLL(drv_path).bench_drive_calls
        drv.restype = C.c_double
        drv.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.c_int, C.c_int]
        n_calls = len(keep)
        p_arr = (C.c_void_p * n_calls)(*[C.addressof(k_[3]) for k_ in keep])
        d_arr = (C.c_void_p * n_calls)(*[C.addressof(k_[2]) for k_ in keep])
        can_p, mm_p = C.cast(can, C.c_void_p), C.cast(mm, C.c_void_p)

    def token(n_tok=1):
        if drv is not None:
            sec_ = drv(can_p, mm_p, p_arr, d_arr, n_calls, n_tok)
            if sec_ < 0:
                raise RuntimeError("lamm_can_mul_mat refused the bench shape")
            return sec_
        t_ = time.perf_counter()
        for _ in range(n_tok):
            for p, d in calls:
                if not can(p, d):
                    raise RuntimeError("lamm_can_mul_mat refused the bench shape")
                mm(p, d)
        return time.perf_counter() - t_
    token()          # uploads the weights
    token()
    torch.cuda.synchronize()
    # the leg's results are checked, not just timed: a sample of the calls against the device-resident entry point
    # on the same weights and the same activation row (same kernels underneath: must agree to f32 noise)
    checked = 0
    for ci in sorted(set([0, 3, 6, len(keep) // 2, len(keep) - 2, len(keep) - 1])):
        name, m, k = ops[ci]
        if not can(*calls[ci]):
            raise RuntimeError("lamm_can_mul_mat refused the bench shape")
        mm(*calls[ci])
        got = ch[m][0, :weights[ci].shape[0]].copy()
        want = L.mul_mat(t, weights[ci] if t != 0 else weights[ci].view(torch.float32),
                         torch.from_numpy(xh[x_of[name]]).to(dev)).cpu().numpy()[0]
        tol = 1e-5 * np.sqrt(k) * float(np.abs(want).max())
        if not np.abs(got - want).max() <= tol:
            raise SystemExit("bench e2e: operator result of call %d (%s) differs from the device path: %g > %g"
                             % (ci, name, float(np.abs(got - want).max()), tol))
        checked += 1
    torch.cuda.synchronize()
    if dist:
        dist.barrier()
    prof = (C.c_double * 8)()
    lib.lammcu_host_profile(prof, 1)
    sec = token(steps)
    lib.lammcu_host_quiesce()                 # outside the timed region: every call above was synchronous
    torch.cuda.synchronize()
    lib.lammcu_host_profile(prof, 1)
    calls = max(prof[5], 1.0)
    phases = {k: round(prof[i] / calls, 2) for i, k in enumerate(
        ("x_memcpy_us", "h2d_enqueue_us", "kernel_enqueue_us", "wait_us", "c_memcpy_us"))}
    phases["total_per_call_us"] = round(sec * 1e6 / calls, 2)
    if dist:
        tt = torch.tensor([sec], device=dev)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        sec = float(tt.item())
    # a row that equals the previous call's is not sent again (host mode 1): count what crossed PCIe
    reused = prof[6] / max(steps, 1)
    h2d_all = sum(k * 4 for _, m, k in ops)
    same_as_prev = [i for i in range(1, len(ops)) if x_of[ops[i][0]] == x_of[ops[i - 1][0]]]
    assert all(ops[i][2] == H for i in same_as_prev)      # every repeated row of this chain has K = H
    h2d = int(h2d_all - reused * H * 4)                   # (the first job of a new server instance always sends its row)
    phases["x_rows_reused_per_step"] = round(reused, 2)
    # calls answered by a job posted ahead of them (sibling speculation: the weight and the row were the predicted ones)
    phases["calls_answered_by_speculative_job_per_step"] = round(prof[7] / max(steps, 1), 2)
    d2h = sum((shard(m)[1] - shard(m)[0]) * 4 for _, m, k in ops)
    return {"value": steps / sec, "unit": "tok/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
            "steps": steps, "ms_per_step": sec / steps * 1e3, "per_call_phases": phases,
            "host_mode": int(lib.lammcu_get_host_mode()), "caller": "C loop (tools/bench_driver.c)" if drv else "python ctypes loop",
            "api": "lamm_can_mul_mat + lamm_mul_mat(ggml_compute_params*, ggml_tensor*) on host tensors, "
                   "one synchronous call per matmul (225 per token); host wall clock; host mode %d (%s)"
                   % (int(lib.lammcu_get_host_mode()),
                      "jobs posted to the resident decode server: X pulled from pinned host memory by the kernel, "
                      "C written back as tagged pairs the caller polls; wq/wk/wv and gate/up get the same src1 "
                      "tensor as in ggml's graph, and a row bit-identical to the previous call's is sent once; the jobs "
                      "of the weights that followed a weight with the same row on the previous token are posted behind "
                      "its job (LAMM_SERVER_SPEC), and a call is answered from such a job only if it is for that weight "
                      "and its row memcmp-equals the job's"
                      if lib.lammcu_get_host_mode() == 1 else

                      "copy-engine H2D + one launch + stream synchronize per call"),
            "distinct_host_weight_buffers": len(host_w), "results_checked_against_device_path": checked,
            "note": "weights uploaded once at first use (registry keyed by src0->data, leaf tensors only)"}


# ---- per-format / per-shape sweep (not the driver's line; writes a table for profiles/) ----
def run_suite(args):
    """GEMV sweep: every format x LLaMA shape at N=args.n, a ring of distinct weight buffers
    >= 512 MB replayed from one CUDA graph (L2-defeating).  Prints one JSON line
    per case: us/launch, GB/s (algorithmic bytes of the whole op incl. the activation quantize when
    --fused), fraction of the measured HBM peak."""
    import torch
    import lamm_b200 as L
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    L.init(0)
    peak, _ = peaks()
    gen = torch.Generator(device=dev)
    gen.manual_seed(7)
    fmts = args.fmts.split(",") if args.fmts else ["q4_0", "q4_1", "q5_0", "q5_1", "q8_0", "q2_k", "f32"]
    shapes = [(4096, 4096), (11008, 4096), (4096, 11008), (32000, 4096), (5120, 5120), (13824, 5120),
              (6656, 6656), (17920, 6656), (6656, 17920)]
    if args.shapes:
        shapes = [tuple(int(v) for v in s_.split("x")) for s_ in args.shapes.split(",")]
    n = args.n
    out = []
    for fmt in fmts:
        t = TYPE_IDS[fmt]
        vt = VEC_DOT[t]
        for (m, k) in shapes:
            if k % BLCK[t]:
                continue
            wb = m * w_row_bytes(t, k)
            ring = max(2, min(96, (args.ring_mb * (1 << 20) + wb - 1) // wb))
            ws = [synth_weight(torch, t, m, k, gen, dev) for _ in range(ring)]
            x = torch.randn((n, k), generator=gen, device=dev)
            xq = L.quantize_act(x, vt) if t != 0 else x
            c = torch.zeros((n, m), device=dev)
            blck, size = ACT_SIZE[vt]
            xq_ws = torch.zeros((n * (k // blck) * size + 64,), dtype=torch.uint8, device=dev)

            def body():
                for w in ws:
                    wv = w if t != 0 else w.view(torch.float32)
                    if args.fused:
                        L.mul_mat(t, wv, x, out=c, ws=xq_ws)
                    else:
                        L.mul_mat_q(t, wv, xq, k, out=c)
            s = torch.cuda.Stream()
            with torch.cuda.stream(s):
                body()
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                body()
            for _ in range(3):
                g.replay()
            torch.cuda.synchronize()
            reps = max(3, int(2000 / ring))
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(reps):
                g.replay()
            e1.record()
            torch.cuda.synchronize()
            us = e0.elapsed_time(e1) * 1e3 / (reps * ring)
            by = gemv_bytes(t, m, k, n) if args.fused else gemv_kernel_bytes(t, m, k, n)
            rec = {"fmt": fmt, "M": m, "K": k, "N": n, "fused_quantize": bool(args.fused), "ring": ring,
                   "us": round(us, 3), "GBps": round(by / us / 1e3, 1), "frac_hbm": round(by / us / 1e3 / peak, 4),
                   "gops": round(2.0 * m * n * k / us / 1e3, 1)}
            print(json.dumps(rec), flush=True)
            out.append(rec)
            del ws, g
            torch.cuda.empty_cache()
    return 0


def run_suite_gemm(args):
    """Prefill sweep: format x LLaMA shape at N=args.n through lammcu_mul_mat_dev (activation quantize +
    operand prep + tcgen05 GEMM), ring of distinct weights, CUDA graph; also the GEMM kernel alone
    (pre-quantized activations still include the prep launch).  TOPS = 2*M*N*K / time."""
    import torch
    import lamm_b200 as L
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    L.init(0)
    pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
    bf16 = json.load(open(pk))["bf16_tflops"] if os.path.exists(pk) else 1590.0
    gen = torch.Generator(device=dev)
    gen.manual_seed(7)
    fmts = args.fmts.split(",") if args.fmts else ["q4_0", "q4_1", "q5_0", "q5_1", "q8_0"]
    shapes = [(4096, 4096), (11008, 4096), (4096, 11008), (5120, 5120), (13824, 5120), (5120, 13824)]
    if args.shapes:
        shapes = [tuple(int(v) for v in s_.split("x")) for s_ in args.shapes.split(",")]
    n = args.n if args.n > 1 else 512
    for fmt in fmts:
        t = TYPE_IDS[fmt]
        for (m, k) in shapes:
            wb = m * w_row_bytes(t, k)
            ring = max(2, min(16, (args.ring_mb * (1 << 20) + wb - 1) // wb))
            ws = [synth_weight(torch, t, m, k, gen, dev) for _ in range(ring)]
            x = torch.randn((n, k), generator=gen, device=dev)
            c = torch.zeros((n, m), device=dev)

            def body():
                for w in ws:
                    L.mul_mat(t, w, x, out=c)
            s = torch.cuda.Stream()
            with torch.cuda.stream(s):
                body()
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                body()
            for _ in range(2):
                g.replay()
            torch.cuda.synchronize()
            reps = max(2, int(200 / ring))
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(reps):
                g.replay()
            e1.record()
            torch.cuda.synchronize()
            us = e0.elapsed_time(e1) * 1e3 / (reps * ring)
            tops = 2.0 * m * n * k / us / 1e6
            print(json.dumps({"fmt": fmt, "M": m, "K": k, "N": n, "ring": ring, "us": round(us, 2),
                              "TOPS": round(tops, 1), "frac_f16_measured": round(tops / bf16, 4),
                              "frac_int8_nominal_4500": round(tops / 4500.0, 4)}), flush=True)
            del ws, g
            torch.cuda.empty_cache()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="cuda", choices=["cuda", "reference"])
    ap.add_argument("--fmt", default="q4_1", choices=sorted(f for f, t in TYPE_IDS.items() if t == 0 or t in _QVAR))
    ap.add_argument("--model", default="7b", choices=sorted(MODELS), help="chain shapes (the metric is quoted on 7b)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-gemm", action="store_true", help="skip the prefill GEMM block of the line")
    ap.add_argument("--no-extras", action="store_true", help="skip the 30B Q5_0 chain / 13B sharded prefill keys")
    ap.add_argument("--sustain-s", type=float, default=2.0, help="seconds of back-to-back replay for the sustained value (0: skip)")
    ap.add_argument("--no-e2e", action="store_true",
                    help="skip the e2e leg (profiler runs only: ncu replays kernels, the resident server talks to the host)")
    ap.add_argument("--host-mode", type=int, default=None, choices=[0, 1],
                    help="e2e leg: completion mode of the host path (lammcu_set_host_mode; default: the library's)")
    ap.add_argument("--no-batch", dest="batch", action="store_false",
                    help="N=1: one launch per matmul (225/token) instead of one per sibling group (129/token)")
    ap.add_argument("--chain", default="plain", choices=["plain", "persistent", "ll"],
                    help="N=1: plain (default) = one PDL-chained launch per sibling group, plain f32 vectors; persistent = "
                         "the token's matmuls as ONE launch (lammcu_chain_*; measured 13 %% slower: DESIGN d1); "
                         "ll = LL {value, tag} vectors")
    ap.add_argument("--collective", default="fused", choices=["fused", "nccl"],
                    help="N>1: how C slices reach every rank (peer-store epilogue, or ncclAllGather per matmul)")
    ap.add_argument("--suite", default=None, choices=[None, "gemv", "gemm"], help="per-format sweep instead of the bench line")
    ap.add_argument("--fmts", default=None)
    ap.add_argument("--shapes", default=None, help="MxK,MxK,...")
    ap.add_argument("--n", type=int, default=1)
    ap.add_argument("--ring-mb", type=int, default=512)
    ap.add_argument("--fused", action="store_true", help="suite: include the activation quantize launch")
    args = ap.parse_args()
    set_model(args.model)
    if args.suite == "gemm":
        return run_suite_gemm(args)
    if args.suite:
        return run_suite(args)
    if args.impl == "reference":
        return run_reference(args)
    return run_cuda(args)


if __name__ == "__main__":
    sys.exit(main())
