"""Small shapes through every kernel family, checked against the oracle.  Written for compute-sanitizer; that tool
turned out to be closed on this pool, so it serves as a fast hang / parity smoke before the full GPU suite:

    python tools/sanitize_small.py [type,type,...]

Decode (fused quantize, K-split), the N = 2..8 GEMV, the prefill GEMMs (both modes), and the host operator through
the resident decode server and through launch-per-call."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402


def main():
    import torch
    import lamm_b200 as L
    from oracle_lib import Oracle, W_TYPES, KQ_TYPES, F32
    torch.cuda.set_device(0)
    L.init(0)
    o = Oracle()
    rng = np.random.default_rng(0)
    bad = 0
    types = [t for t in W_TYPES + KQ_TYPES]
    if len(sys.argv) > 1:
        types = [int(v) for v in sys.argv[1].split(",")]
    for t in types:
        for (m, n, k) in [(300, 1, 1024), (70, 1, 5120), (130, 5, 512), (200, 40, 512), (129, 130, 256), (260, 300, 2048)]:
            w = rng.standard_normal((m, k)).astype(np.float32) if t == F32 else \
                o.quantize_weights(t, rng.standard_normal((m, k)).astype(np.float32), seed=m)
            x = rng.standard_normal((n, k)).astype(np.float32)
            want = o.mul_mat(t, w if t != F32 else w.view(np.uint8).reshape(m, -1), x)
            for exact in ((0, 1) if n > 8 else (0,)):
                L.lib().lammcu_set_gemm_exact(exact)
                got = L.mul_mat(t, torch.from_numpy(w).cuda(), torch.from_numpy(x).cuda()).cpu().numpy()
                L.lib().lammcu_set_gemm_exact(0)
                ok = np.abs(got - want).max() <= 1e-5 * np.sqrt(k) * np.abs(want).max()
                bad += 0 if ok else 1
                print(f"type {t} {m}x{n}x{k} exact={exact}: {'ok' if ok else 'FAIL'}", flush=True)
        if t != F32:                                     # host operator, decode: resident server and launch-per-call
            m, k = 200, 1024
            w = o.quantize_weights(t, rng.standard_normal((m, k)).astype(np.float32), seed=1)
            x = rng.standard_normal((1, k)).astype(np.float32)
            want = o.mul_mat(t, w, x)
            for mode in (1, 0):
                L.lib().lammcu_set_host_mode(mode)
                can, got = L.lamm_mul_mat_host(t, w, x)
                ok = can and np.abs(got - want).max() <= 1e-5 * np.sqrt(k) * np.abs(want).max()
                bad += 0 if ok else 1
                print(f"type {t} operator host mode {mode}: {'ok' if ok else 'FAIL'}", flush=True)
            L.lib().lammcu_set_host_mode(1)
            L.lib().lammcu_weight_forget_all()
    L.lib().lammcu_host_quiesce()
    torch.cuda.synchronize()
    print("sanitize_small:", "FAILED" if bad else "all ok")
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
