"""Generates tests/golden/search_small.npz: small known-answer cases of the DEFINED search.

Not pinned by the reference (it has no search to pin) — these are regression vectors of
the oracle's exhaustive defined search (oracle/flat_oracle.c: b2r_oracle_search), so that the oracle
restatement itself cannot drift silently and the GPU box can check the CUDA path against committed data.

    python tests/golden/make_search_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import flat_oracle as fo  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "search_small.npz")


def main():
    rng = np.random.default_rng(7)
    out = {}
    specs = [  # n, d, nq, k, metric, normalise, duplicates
        (1000, 96, 7, 5, 0, True, False),
        (777, 100, 9, 10, 1, False, False),
        (512, 64, 4, 16, 0, True, True),
        (300, 384, 3, 4, 1, True, True),
        (20, 32, 5, 32, 0, False, False),   # k > n: padding
    ]
    for ci, (n, d, nq, k, metric, norm, dup) in enumerate(specs):
        x = rng.standard_normal((n, d)).astype(np.float32)
        q = rng.standard_normal((nq, d)).astype(np.float32)
        if dup:
            x[n // 2:] = x[: n - n // 2]          # every row appears twice: ties -> lower id
            q[0] = x[3]
        if norm:
            x, q = fo.normalize(x), fo.normalize(q)
        D, I = fo.search_defined(x, q, k, metric)
        out[f"c{ci}_x"], out[f"c{ci}_q"], out[f"c{ci}_D"], out[f"c{ci}_I"] = x, q, D, I
        out[f"c{ci}_meta"] = np.array([k, metric])
    out["ncases"] = np.array(len(specs))
    np.savez_compressed(OUT, **out)
    print("wrote", OUT)


if __name__ == "__main__":
    main()
