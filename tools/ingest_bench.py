"""Kernel (a) throughput: ingest (cast/copy + row norms) and in-place normalisation.  tools/ingest_bench.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rag_tl_domainllm_optimizer_b200 as b2r  # noqa: E402

n, d = 1_000_000, 768
x = torch.randn((n, d), device="cuda")
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for storage, normalize in (("f32", False), ("bf16", False), ("f32", True), ("bf16", True)):
    best = 1e9
    for _ in range(4):
        idx = b2r.IndexFlatIP(d, storage=storage, normalize=normalize)
        idx.reserve(n)
        e0.record(); idx.add(x); e1.record(); e1.synchronize()
        best = min(best, e0.elapsed_time(e1))
        del idx
    es = 4 if storage == "f32" else 2
    print(f"ingest storage={storage} normalize={normalize}: {best:.3f} ms  {n * d * (4 + es) / best / 1e6:.0f} GB/s algorithmic")
best = 1e9
for _ in range(4):
    y = x.clone()
    e0.record(); b2r.normalize_L2(y); e1.record(); e1.synchronize()
    best = min(best, e0.elapsed_time(e1))
print(f"normalize_L2 in place: {best:.3f} ms  {n * d * 8 / best / 1e6:.0f} GB/s algorithmic (read + write once)")
