#!/usr/bin/env python
"""bench.py — flat top-k search throughput (queries/sec) on B200, one JSON line.

Workload (BASELINE.json configs[1]): 1M x 768 corpus (all-mpnet-base dim), k=10, unit-norm synthetic
embeddings; headline query batch 1024, exact fp32 mode; `--sweep` adds the other batch sizes / the bf16
mode.  A "step" is one search() of the whole batch over the whole corpus.  With N > 1 ranks (torchrun) the
SAME corpus is row-sharded over the ranks (strong scaling) and every search ends with one all-gather of the
packed shard-local top-k plus the merge kernel.

  value     : queries/sec with queries and outputs already resident in HBM (CUDA events, max over ranks)
  e2e       : the same through the public API with HOST buffers (H2D of the queries and D2H of D/I inside
              the timed region, every step)
  roofline  : the scoring kernel's algorithmic bytes (N_local*d*s) or flops (2*B*N_local*d) per search over
              its own CUDA-event duration inside the timed region, against MEASURED_PEAKS.json
  cpu_baseline : the oracle's FAISS-CPU-like fp32 search ("port") on this box's host cores, bounded sample

`--impl reference` times that CPU path alone (rank 0 only) and prints the same line with "impl": "reference".
"""
import argparse
import json
import os
import statistics
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

N_ROWS, DIM, K = 1_000_000, 768, 10
# dram__bytes_read.sum + dram__bytes_write.sum of the main scoring launch, one `ncu --set full` capture each
# (profiles/r01_ncu_full_*.csv); keyed by (mode, batch, rows, n_gpus)
NCU_TRAFFIC = {("bf16", 32, 1_000_000, 1): 1.5173e9, ("exact", 32, 1_000_000, 1): 3.0329e9,
               ("bf16", 1024, 1_000_000, 1): 1.5507e9, ("exact", 1024, 1_000_000, 1): 3.1153e9}
FALLBACK_PEAKS = {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return d, "measured"
    return dict(FALLBACK_PEAKS), "fallback"


class ClockSampler(threading.Thread):
    """Samples SM clock / throttle reasons through NVML while the timed region runs."""

    def __init__(self, dev: int, period: float = 0.1):
        super().__init__(daemon=True)
        self.dev, self.period, self.samples, self.reasons, self.max_mhz = dev, period, [], set(), None
        self._stop_evt = threading.Event()
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(dev)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            self.ok = False

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
            getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonHwPowerBrakeSlowdown", 0x80): "hw_power_brake",
        }
        while not self._stop_evt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    mask = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(self.period)

    def stop(self):
        self._stop_evt.set()
        if self.is_alive():
            self.join(2)
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": []}
        return {"sm_mhz": statistics.median(self.samples), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}


def gen_rows_gpu(torch, seed, row0, nrows, d, device):
    """Unit-norm synthetic rows, reproducible per 65536-row block whatever the sharding."""
    out = torch.empty((nrows, d), dtype=torch.float32, device=device)
    blk = 65536
    b0 = row0 // blk * blk
    while b0 < row0 + nrows:
        g = torch.Generator(device=device)
        g.manual_seed(seed * 1_000_003 + b0 // blk)
        rows = torch.randn((blk, d), generator=g, dtype=torch.float32, device=device)
        lo, hi = max(b0, row0), min(b0 + blk, row0 + nrows)
        out[lo - row0:hi - row0] = rows[lo - b0:hi - b0]
        b0 += blk
    return out


def cpu_search_sample(fo, x_host, q_host, k, metric, nq_sample):
    qs = q_host[:nq_sample]
    t0 = time.perf_counter()
    fo.search_fp32_blas(x_host, qs, k, metric)
    return time.perf_counter() - t0


def run_reference(args, rank, world):
    """The reference arm: the CPU restatement of FAISS-flat search on this box's host cores."""
    if rank != 0:
        return
    import torch
    from oracle import flat_oracle as fo
    fo.build()
    cores = len(os.sched_getaffinity(0))
    torch.set_num_threads(cores)
    rng = np.random.default_rng(0)
    x = rng.standard_normal((args.rows, DIM), dtype=np.float32)
    x /= np.linalg.norm(x, axis=1, keepdims=True)
    q = rng.standard_normal((args.batch, DIM), dtype=np.float32)
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    nqs = min(args.batch, args.cpu_queries)          # 256 queries: ~9 s per step on 16 cores
    for _ in range(min(args.warmup, 1)):
        cpu_search_sample(fo, x, q, args.k, 0, nqs)
    t = [cpu_search_sample(fo, x, q, args.k, 0, nqs) for _ in range(args.steps)]
    sec = sum(t) / len(t)
    qps = nqs / sec
    sample = f"{nqs} of the {args.batch} queries per step over the full {args.rows}x{DIM} corpus"
    line = {
        "impl": "reference", "metric": "flat top-k search queries/sec", "value": qps, "unit": "queries/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec * 1e3,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, world),
        "cpu_baseline": {"value": qps, "unit": "queries/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": qps, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "CPU restatement of FAISS-flat (NumPy/OpenBLAS sgemm + C heap), not FAISS: the reference pins no "
                "FAISS version and FAISS is not installable here (BASELINE.md §3)",
    }
    print(json.dumps(line), flush=True)


def workload_config(args, world):
    return {"workload": f"{'configs[1]: ' if (args.rows, DIM) == (N_ROWS, 768) else ''}{args.rows}x{DIM} corpus, query batch {args.batch}, k={args.k}, "
                        f"{'exact fp32' if args.mode == 'exact' else 'bf16'} inner-product flat search",
            "rows": args.rows, "dim": DIM, "batch": args.batch, "k": args.k,
            "mode": args.mode + ("+bf16-shadow prefilter" if getattr(args, "shadow", False) and args.mode == "exact" else ""),
            "sharding": f"rows/{world}" if world > 1 else "none",
            "l2": "inputs larger than L2: corpus shard re-read from HBM every step (>= 126 MB)"}


def main():
    global DIM
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=1024)
    ap.add_argument("--k", type=int, default=K)
    ap.add_argument("--rows", type=int, default=N_ROWS)
    ap.add_argument("--dim", type=int, default=DIM, help="embedding dimension (other BASELINE configs)")
    ap.add_argument("--mode", default="exact", choices=["exact", "bf16"])
    ap.add_argument("--engine", type=int, default=0)
    ap.add_argument("--shadow", action="store_true", help="exact mode: score a bf16 shadow copy, re-score fp32 rows")
    ap.add_argument("--sweep", default="1,8,32,128,1024", help="comma list of extra batch sizes ('' = none)")
    ap.add_argument("--cpu-queries", type=int, default=256)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--check-queries", type=int, default=32)
    args = ap.parse_args()
    DIM = args.dim
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else max(args.warmup, 1)

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch
    import torch.distributed as dist

    import rag_tl_domainllm_optimizer_b200 as b2r

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the library has no CPU path")
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=device)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(v: float) -> float:
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    peaks, peak_src = load_peaks()
    storage = "f32" if args.mode == "exact" else "bf16"
    esize = 4 if storage == "f32" else 2
    esize_score = 2 if (args.shadow and storage == "f32") else esize   # bytes per element the scoring pass reads

    # ---- corpus shard + queries -------------------------------------------------------------
    lo, hi = b2r.shard_bounds(args.rows, world, rank)
    sidx = b2r.ShardedIndexFlat(DIM, b2r.METRIC_INNER_PRODUCT, storage=storage, device=local_rank,
                                shadow=args.shadow and storage == "f32")
    sidx.local.set_engine(args.engine)
    sidx.local.reserve(hi - lo)
    chunk = 131072
    ing0, ing1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ingest_ms = 0.0
    for r0 in range(lo, hi, chunk):
        rows = gen_rows_gpu(torch, 0, r0, min(chunk, hi - r0), DIM, device)
        b2r.normalize_L2(rows)                      # kernel (a), in place, as faiss.normalize_L2 is used
        ing0.record()
        sidx.add_local(rows, lo, args.rows)         # kernel (a) again: cast / copy into the index, row norms
        ing1.record()
        ing1.synchronize()
        ingest_ms += ing0.elapsed_time(ing1)
        del rows
    torch.cuda.synchronize()
    ingest_gbs = (hi - lo) * DIM * (4 + esize) / (ingest_ms * 1e-3) / 1e9 if ingest_ms else 0.0

    def make_queries(b):
        q = gen_rows_gpu(torch, 1, 0, b, DIM, device)
        b2r.normalize_L2(q)
        return q

    def timed_device(q, steps, warmup):
        for _ in range(warmup):
            sidx.search(q, args.k)
        sidx.local.score_time(reset=True)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            sidx.search(q, args.k)
        e1.record()
        barrier()
        ms = max_over_ranks(e0.elapsed_time(e1))
        tot, cnt = sidx.local.score_time(reset=True)
        return ms / steps, (tot / cnt if cnt else 0.0)

    def timed_e2e(q_host, steps, warmup):
        for _ in range(warmup):
            sidx.search(q_host, args.k)
        barrier()
        t0 = time.perf_counter()
        for _ in range(steps):
            D, I = sidx.search(q_host, args.k)
            D_h = D.cpu() if hasattr(D, "cpu") else torch.from_numpy(D)     # device -> host read of the step's result
            I_h = I.cpu() if hasattr(I, "cpu") else torch.from_numpy(I)     # (host-buffer callers get it from the C ABI)
        barrier()
        return max_over_ranks((time.perf_counter() - t0) * 1e3) / steps, D_h, I_h

    # ---- headline ----------------------------------------------------------------------------
    q = make_queries(args.batch)
    sampler = ClockSampler(local_rank)
    sampler.start()
    ms_step, score_ms = timed_device(q, args.steps, args.warmup)
    clocks = sampler.stop()
    st = sidx.local.last_stats()
    launches_per_step = st["launches"] + (1 if world > 1 else 0)   # + the merge kernel of the gathered lists
    qps = args.batch / (ms_step * 1e-3)

    q_pinned = torch.empty((args.batch, DIM), dtype=torch.float32, pin_memory=True)
    q_pinned.copy_(q)
    q_host = q_pinned.numpy()
    e2e_ms, D_h, I_h = timed_e2e(q_host, args.steps, args.warmup)
    e2e_qps = args.batch / (e2e_ms * 1e-3)

    # ---- roofline of the dominant (scoring) kernel -------------------------------------------
    n_local = hi - lo
    rows_k = (st["main_rows"] or n_local) if st["engine"] == 2 else n_local    # rows the timed (dominant) launch covers
    bytes_alg = rows_k * DIM * esize_score
    flops_alg = 2.0 * args.batch * rows_k * DIM
    hbm_gbs = bytes_alg / (score_ms * 1e-3) / 1e9 if score_ms else 0.0
    tflops = flops_alg / (score_ms * 1e-3) / 1e12 if score_ms else 0.0
    hbm_frac = hbm_gbs / peaks["hbm_gbs"]
    # the timed region is tens of ms (no power-cap settling): compare with the BURST cuBLAS figure, the harder one;
    # the fraction of the sustained figure is reported beside it
    tc_peak = peaks["bf16_tflops"]
    tc_sustained = peaks.get("bf16_tflops_sustained", tc_peak)
    tc_frac = tflops / tc_peak
    if hbm_frac >= tc_frac:
        roof = {"bound": "hbm", "achieved": hbm_gbs, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": hbm_frac}
    else:
        roof = {"bound": "tensor", "achieved": tflops, "peak": tc_peak, "unit": "TFLOP/s", "frac": tc_frac,
                "frac_of_sustained_peak": tflops / tc_sustained}
    if storage == "f32" and st["engine"] in (2, 3):
        # exact mode runs kind::tf32 MMAs, whose dense rate is half the bf16 rate (1.1 vs 2.25 PFLOP/s nominal);
        # MEASURED_PEAKS.json has no tf32 figure, so the fraction against half the measured bf16 peak is added
        roof["frac_of_tf32_peak_derived"] = tflops / (tc_peak / 2.0)
    roof.update({"traffic": NCU_TRAFFIC.get((args.mode, args.batch, args.rows, world)),
                 "traffic_source": "profiles/README.md (ncu --set full, dram read+write of the same launch)",
                 "kernel": "score_umma_kernel (main launch)" if st["engine"] == 2 else "score_fma_kernel (all passes)",
                 "rows_per_launch": rows_k, "kernel_ms": score_ms, "peak_source": f"{peak_src} (bf16 burst {tc_peak} / sustained {tc_sustained} TFLOP/s; hbm copy)",
                 "algorithmic": {"bytes_per_launch_set": bytes_alg, "flops_per_launch_set": flops_alg}})

    # ---- parity spot-check and CPU baseline (rank 0, host cores) ------------------------------
    parity, cpu_base = None, None
    if rank == 0 and world == 1 and not args.no_cpu:
        from oracle import flat_oracle as fo
        fo.build()
        x_host = torch.empty((args.rows, DIM), dtype=torch.float32)
        for r0 in range(0, args.rows, chunk):
            rows = gen_rows_gpu(torch, 0, r0, min(chunk, args.rows - r0), DIM, device)
            b2r.normalize_L2(rows)
            x_host[r0:r0 + rows.shape[0]] = rows.cpu()
        xh = x_host.numpy()
        nc = min(args.check_queries, args.batch)
        xo, qo = (fo.round_bf16(xh), fo.round_bf16(q_host[:nc])) if storage == "bf16" else (xh, q_host[:nc])
        D0, I0 = fo.search_ref(xo, qo, args.k, 0)
        ids_equal = bool(np.array_equal(I_h.numpy()[:nc], I0))
        scores_equal = bool(np.array_equal(D_h.numpy()[:nc].view(np.uint32), D0.view(np.uint32)))
        parity = {"checked_queries": nc, "ids_identical": ids_equal, "scores_bit_identical": scores_equal,
                  "unverified_queries": st["unverified"]}
        if storage == "bf16":
            Df, If = fo.search_ref(xh, q_host[:nc], args.k, 0)
            parity["recall_at_k_vs_fp32"] = float(np.mean([len(set(a) & set(b)) / args.k
                                                           for a, b in zip(I_h.numpy()[:nc], If)]))
        cores = len(os.sched_getaffinity(0))
        nqs = min(args.batch, args.cpu_queries)
        cpu_search_sample(fo, xh, q_host, args.k, 0, min(nqs, 32))      # warm the BLAS threads
        sec = cpu_search_sample(fo, xh, q_host, args.k, 0, nqs)
        cpu_base = {"value": nqs / sec, "unit": "queries/s", "cores": cores, "kind": "port",
                    "sample": f"{nqs} of the {args.batch} queries, one pass over the full {args.rows}x{DIM} corpus "
                              f"({sec:.2f} s); CPU restatement of FAISS-flat, not FAISS"}
        del x_host, xh

    # ---- sweep over the other batch sizes / modes (1 GPU only) --------------------------------
    sweep = []
    if world == 1 and args.sweep:
        for b in [int(v) for v in args.sweep.split(",") if v]:
            if b == args.batch:
                sweep.append({"batch": b, "mode": args.mode, "qps": qps, "score_ms": score_ms,
                              "hbm_frac": hbm_frac, "tensor_frac": tc_frac})
                continue
            qb = make_queries(b)
            steps = max(5, min(args.steps, 50))
            ms_b, sc_b = timed_device(qb, steps, 3)
            sweep.append({"batch": b, "mode": args.mode, "qps": b / (ms_b * 1e-3), "score_ms": sc_b,
                          "hbm_frac": (bytes_alg / (sc_b * 1e-3) / 1e9) / peaks["hbm_gbs"] if sc_b else 0.0,
                          "tensor_frac": (2.0 * b * rows_k * DIM / (sc_b * 1e-3) / 1e12) / tc_peak if sc_b else 0.0})

    if rank == 0:
        line = {
            "metric": "flat top-k search queries/sec", "value": qps, "unit": "queries/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None,
            "dtype": {2: "tf32", 3: "tf32", 4: "bf16"}.get(st["engine"], "f32") if storage == "f32" else "bf16",
            "data": "synthetic", "config": workload_config(args, world), "clocks": clocks,
            "e2e": {"value": e2e_qps, "unit": "queries/s", "ms_per_step": e2e_ms,
                    "h2d_bytes_per_step": args.batch * DIM * 4, "d2h_bytes_per_step": args.batch * args.k * 12},
            "gpu_launches": launches_per_step * args.steps,
            "roofline": roof, "cpu_baseline": cpu_base, "parity": parity, "sweep": sweep,
            "engine": {1: "fma", 2: "umma", 3: "umma-3xtf32", 4: "umma-bf16-shadow"}.get(st["engine"], st["engine"]), "kprime": st["kprime"],
            "ingest": {"kernel": "ingest_rows_kernel", "gbs": ingest_gbs, "frac_of_hbm_peak": ingest_gbs / peaks["hbm_gbs"],
                       "bytes_per_row": DIM * (4 + esize)},
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
