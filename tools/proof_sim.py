"""CPU simulation of the exact-mode completeness proof on synthetic distributions (no GPU needed).

Emulates the 1xTF32 scoring pass (operands truncated to 10 mantissa bits) with NumPy, keeps the best K'
approximate scores per query and evaluates the proof  fl32(a_min + E) < s_k  of DESIGN.md §3.4 for several
(K', E) pairs.  Used to size K' and to predict the share of queries that take the second (range) pass on
the distributions bench.py --dist generates.
"""
import sys
import time

import numpy as np


def trunc_tf32(a):
    return (a.view(np.uint32) & np.uint32(0xFFFFE000)).view(np.float32)


def round_bf16(a):
    u = a.view(np.uint32).astype(np.uint64)
    u = (u + 0x7FFF + ((u >> 16) & 1)) & 0xFFFF0000
    return u.astype(np.uint32).view(np.float32)


def gen(dist, n, d, nq, seed=0):
    rng = np.random.default_rng(seed)
    if dist == "gauss":
        x = rng.standard_normal((n, d), dtype=np.float32)
        q = rng.standard_normal((nq, d), dtype=np.float32)
    else:
        nc = 1024
        c = rng.standard_normal((nc, d), dtype=np.float32)
        c /= np.linalg.norm(c, axis=1, keepdims=True)
        sig = np.float32(0.3 / np.sqrt(d))
        x = np.empty((n, d), dtype=np.float32)
        for r0 in range(0, n, 65536):
            m = min(65536, n - r0)
            x[r0:r0 + m] = c[rng.integers(0, nc, m)] + sig * rng.standard_normal((m, d), dtype=np.float32)
        src = rng.integers(0, n, nq)
        q = x[src] + sig * rng.standard_normal((nq, d), dtype=np.float32)
        if dist == "dup":
            ndup = n // 100
            x[rng.integers(0, n, ndup)] = x[rng.integers(0, n, ndup)]
    x /= np.linalg.norm(x, axis=1, keepdims=True)
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    return x, q


def main():
    dist = sys.argv[1] if len(sys.argv) > 1 else "mixture"
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 1_000_000
    nq = int(sys.argv[3]) if len(sys.argv) > 3 else 256
    d, k, keep = 768, 10, 512
    t0 = time.time()
    x, q = gen(dist, n, d, nq)
    print(f"{dist}: generated {n}x{d} in {time.time()-t0:.1f}s", flush=True)
    modes = {"tf32": (trunc_tf32, (d / 4 + 8) * 1.1921e-7 + 1.96e-3), "bf16": (round_bf16, (d / 8 + 8) * 1.1921e-7 + 3.92e-3)}
    for name, (fn, errfac) in modes.items():
        qa = fn(q)
        best_a = np.full((nq, 0), 0, np.float32); best_e = np.full((nq, 0), 0, np.float32)
        for r0 in range(0, n, 131072):
            xb = x[r0:r0 + 131072]
            sa = qa @ fn(xb).T
            se = q @ xb.T
            sa = np.concatenate([best_a, sa], 1); se = np.concatenate([best_e, se], 1)
            part = np.argpartition(-sa, keep - 1, axis=1)[:, :keep]
            best_a = np.take_along_axis(sa, part, 1); best_e = np.take_along_axis(se, part, 1)
        order = np.argsort(-best_a, axis=1)
        best_a = np.take_along_axis(best_a, order, 1); best_e = np.take_along_axis(best_e, order, 1)
        err = np.abs(best_a - best_e).max()
        print(f"[{name}] max |approx - exact| over kept = {err:.2e} (bound {errfac:.2e})")
        for kp in (16, 32, 64, 128, 256):
            sk = -np.sort(-best_e[:, :kp], axis=1)[:, k - 1]
            amin = best_a[:, kp - 1]
            for E, lab in ((errfac, "worst-case"), (errfac / 4, "E/4"), (errfac / 16, "E/16")):
                fail = np.mean(amin + E >= sk)
                print(f"  K'={kp:4d} {lab:10s}: unproven {100*fail:6.2f}%")
        # range pass: how many rows have approx >= s_k - E  (s_k from K'=32)
        sk = -np.sort(-best_e[:, :32], axis=1)[:, k - 1]
        cnt = (best_a >= (sk - errfac)[:, None]).sum(1)
        print(f"  range pass 1x ({name}): rows with approx >= s_k - E: median {np.median(cnt):.0f} p90 {np.percentile(cnt,90):.0f} max {cnt.max()} (cap {keep})")
        cnt3 = (best_e >= (sk - 3e-6)[:, None]).sum(1)
        print(f"  range pass 3xTF32: rows within 3e-6 of s_k or better: max {cnt3.max()}")


if __name__ == "__main__":
    main()
