"""GPU test of the allocation-free reward-cosine entry point, `b2r_cosine_rows_noalloc`.

That entry point is for callers that cannot allocate device memory at call time: the reward step of a PPO loop
running beside a policy whose framework has pooled all free HBM (a raw cudaMalloc there may fail and always stalls
the device), or a serving loop that forbids allocation after warm-up.  It must give exactly what
`b2r_cosine_rows` gives.  It keeps its scratch in one fixed device array that every call reuses, so the test makes
calls in a row with different inputs and sizes (smaller, then larger than before): stale maxima from an earlier
call must not leak into a later one.  Outputs are also checked against the CPU oracle directly.

The file name sorts last on purpose: a `-x` run has already run every other test when it gets here.  The test also
passes alone in a fresh process (the entry point resolves its scratch address on first use).
"""
import ctypes

import numpy as np
import pytest

from rag_tl_domainllm_optimizer_b200 import _lib

pytestmark = pytest.mark.gpu


def _ref(L, resp, docs, torch):
    """The existing entry point, on the same device buffers."""
    n, m, d = docs.shape
    cos = torch.empty((n, m), dtype=torch.float32, device="cuda")
    mx = torch.empty((n,), dtype=torch.float32, device="cuda")
    rc = L.b2r_cosine_rows(n, m, d, resp.data_ptr(), docs.data_ptr(), cos.data_ptr(), mx.data_ptr(), 0, None)
    assert rc == 0, L.b2r_last_error()
    torch.cuda.synchronize()
    return cos.cpu().numpy(), mx.cpu().numpy()


def _noalloc(L, resp, docs, torch):
    n, m, d = docs.shape
    cos = torch.full((n, m), -7.0, dtype=torch.float32, device="cuda")
    mx = torch.full((n,), -7.0, dtype=torch.float32, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    rc = L.b2r_cosine_rows_noalloc(n, m, d, resp.data_ptr(), docs.data_ptr(), cos.data_ptr(), mx.data_ptr(), 0, stream)
    assert rc == 0, L.b2r_last_error()
    torch.cuda.synchronize()
    return cos.cpu().numpy(), mx.cpu().numpy()


def test_noalloc_matches_allocating_entry_point_twice_in_a_row(oracle):
    import torch
    L = _lib.lib()
    g = torch.Generator(device="cuda")
    g.manual_seed(5)
    # call 1: large positive cosines everywhere (docs are scaled copies of the responses plus noise)
    resp1 = torch.randn((300, 768), generator=g, device="cuda")
    docs1 = resp1[:, None, :] * 2.0 + 0.05 * torch.randn((300, 4, 768), generator=g, device="cuda")
    docs1 = docs1.contiguous()
    # call 2: fewer samples, another dimension, negative cosines — a stale maximum from call 1 would win
    resp2 = torch.randn((77, 384), generator=g, device="cuda")
    docs2 = (-resp2[:, None, :] + 0.3 * torch.randn((77, 3, 384), generator=g, device="cuda")).contiguous()
    # call 3: MORE samples than either call before: entries 300.. of the scratch have never been used or cleared
    resp3 = torch.randn((1000, 64), generator=g, device="cuda")
    docs3 = (-resp3[:, None, :] * 0.5 + 0.2 * torch.randn((1000, 2, 64), generator=g, device="cuda")).contiguous()
    for resp, docs in ((resp1, docs1), (resp2, docs2), (resp3, docs3)):
        cos0, mx0 = _ref(L, resp, docs, torch)
        cos1, mx1 = _noalloc(L, resp, docs, torch)
        # same kernels, same arithmetic: the tolerance of the parity tests is bit equality
        assert np.array_equal(cos1.view(np.uint32), cos0.view(np.uint32))
        assert np.array_equal(mx1.view(np.uint32), mx0.view(np.uint32))
        # and directly against the CPU oracle's restatement of the reference's cosine (reference :63-71)
        cos_o, mx_o = oracle.cosine_rows(resp.cpu().numpy(), docs.cpu().numpy())
        assert np.array_equal(cos1.view(np.uint32), cos_o.view(np.uint32))
        assert np.array_equal(mx1.view(np.uint32), mx_o.view(np.uint32))
    assert np.all(mx1 < 0) and np.all(_ref(L, resp1, docs1, torch)[1] > 0.9)


def test_noalloc_leaves_free_memory_untouched():
    """Steady state: the driver's count of free device memory is the same before and after a run of calls (the
    allocating entry point, by contrast, takes and returns a workspace on every call)."""
    import torch
    L = _lib.lib()
    resp = torch.randn((512, 256), device="cuda")
    docs = torch.randn((512, 3, 256), device="cuda")
    cos = torch.empty((512, 3), device="cuda")
    mx = torch.empty((512,), device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    args = (512, 3, 256, resp.data_ptr(), docs.data_ptr(), cos.data_ptr(), mx.data_ptr(), 0, stream)
    assert L.b2r_cosine_rows_noalloc(*args) == 0          # first call on the device: resolves the scratch address
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    for _ in range(20):
        assert L.b2r_cosine_rows_noalloc(*args) == 0
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0


def test_noalloc_limits_and_arguments():
    import torch
    L = _lib.lib()
    x = torch.zeros((4, 8), device="cuda")
    out = torch.zeros((4,), device="cuda")
    too_many = 65536 + 1
    rc = L.b2r_cosine_rows_noalloc(too_many, 1, 8, x.data_ptr(), x.data_ptr(), None, out.data_ptr(), 0, None)
    assert rc == -5 and b"B2R_NOALLOC_MAX_SAMPLES" in L.b2r_last_error()
    host = np.zeros((4, 8), dtype=np.float32)
    rc = L.b2r_cosine_rows_noalloc(4, 1, 8, host.ctypes.data, x.data_ptr(), None, out.data_ptr(), 0, None)
    assert rc == -1 and b"device pointers only" in L.b2r_last_error()
    if torch.cuda.device_count() > 1:                     # a buffer of another GPU is refused, not dereferenced
        rc = L.b2r_cosine_rows_noalloc(4, 1, 8, x.data_ptr(), x.data_ptr(), None, out.data_ptr(), 1, None)
        assert rc == -1 and b"lives on device" in L.b2r_last_error()
    # m == 0: the maximum over no documents is 0.0, as at reference :71
    out.fill_(3.0)
    rc = L.b2r_cosine_rows_noalloc(4, 0, 8, x.data_ptr(), None, None, out.data_ptr(), 0, None)
    assert rc == 0
    torch.cuda.synchronize()
    assert np.all(out.cpu().numpy() == 0.0)
