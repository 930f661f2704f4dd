"""Smallest end-to-end case per engine, for compute-sanitizer (memcheck): tools/sanitize_case.py"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rag_tl_domainllm_optimizer_b200 as b2r  # noqa: E402

rng = np.random.default_rng(0)
x = rng.standard_normal((700, 96)).astype(np.float32)
q = rng.standard_normal((5, 96)).astype(np.float32)
for storage, engine in (("f32", b2r.ENGINE_FMA), ("f32", b2r.ENGINE_UMMA), ("f32", b2r.ENGINE_UMMA3), ("bf16", b2r.ENGINE_UMMA)):
    for metric in (0, 1):
        idx = b2r.IndexFlat(96, metric, storage=storage, normalize=True)
        idx.set_engine(engine)
        idx.add(x)
        D, I = idx.search(q, 7)
        print(storage, engine, metric, I[0, :3], idx.last_stats()["engine"])
cos, mx = b2r.cosine_rows(q, np.stack([x[:15].reshape(5, 3, 96)])[0])
print("cosine", mx[:2])
