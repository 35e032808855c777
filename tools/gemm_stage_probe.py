"""Which weight data did each K stage of the pair GEMM actually multiply?  (diagnostics; this probe found the raw-buffer
release race of round 2, profiles/r2bf)

    python tools/gemm_stage_probe.py            # LAMM_GEMM_PAIR=2 LAMM_GEMM_ATM=1 unless set

Activation row b is an indicator of K block b, so column b of C is the sum of block b of every weight row (Q5_1 with
m = 0: no min term).  A wrong cell names the (stage, unpack half, row) whose tile was wrong; comparing it with the other
columns of the same row tells whose data the tile held instead (the same position one chunk later = the raw chunk was
overwritten by the next TMA load before the row's last ld.shared had executed)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402


def main():
    import torch
    import lamm_b200 as L
    from oracle_lib import Oracle
    torch.cuda.set_device(0)
    L.init(0)
    o = Oracle()
    os.environ.setdefault("LAMM_GEMM_PAIR", "2")
    os.environ.setdefault("LAMM_GEMM_ATM", "1")
    t, m, n, k = 7, 256, 256, 512
    rng = np.random.default_rng(m + n)
    w = o.quantize_weights(t, rng.standard_normal((m, k)).astype(np.float32), seed=m + n).reshape(m, k // 32, 24)
    w[:, :, 2:4] = 0                      # m = 0: no min term
    w = w.reshape(m, -1)
    x = np.zeros((n, k), np.float32)
    nblk = k // 32
    for b in range(nblk):
        x[b, 32 * b:32 * b + 32] = 1.0
    x[nblk:] = np.random.default_rng(1).standard_normal((n - nblk, k)).astype(np.float32)
    want = o.mul_mat(t, w, x)
    total = 0
    for rep in range(4):
        got = L.mul_mat(t, torch.from_numpy(w).cuda(), torch.from_numpy(x).cuda()).cpu().numpy()
        d = np.abs(got[:nblk] - want[:nblk])
        bad = np.argwhere(d > 1e-3 * np.abs(want[:nblk]).max())
        total += len(bad)
        print(f"rep {rep}: {len(bad)} bad (block, row) cells; blocks {sorted(set(bad[:, 0].tolist()))}")
        for (b, r) in bad[:6]:
            match = [bb for bb in range(nblk) if abs(want[bb, r] - got[b, r]) < 1e-3 * max(1, abs(want[bb, r]))]
            print(f"   block {b} (stage {b // 2}, half {b & 1}) row {r}: got {got[b, r]:.4f} want {want[b, r]:.4f}; "
                  f"equals the sum of block(s) {match}")
    sys.exit(1 if total else 0)


if __name__ == "__main__":
    main()
