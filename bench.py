#!/usr/bin/env python
"""bench.py — flat top-k search throughput (queries/sec) on B200, one JSON line.

Workload (BASELINE.json configs[1]): 1M x 768 corpus (all-mpnet-base dim), k=10, unit-norm synthetic
embeddings; headline query batch 1024, exact fp32 mode.  A "step" is one search() of the whole batch over the
whole corpus.  With N > 1 ranks (torchrun) the SAME corpus is row-sharded over the ranks (strong scaling) and
every search ends with one NCCL all-gather of the packed shard-local top-k plus the merge kernel
(b2r_index_search_sharded).

  value     : queries/sec with queries and outputs already resident in HBM (CUDA events, max over ranks)
  e2e       : the same through the public API with HOST buffers (H2D of the queries and D2H of D/I inside
              the timed region, every step)
  roofline  : the scoring kernel's algorithmic bytes (N_local*d*s) or flops (2*B*N_local*d) per search over
              its own CUDA-event duration inside the timed region, against MEASURED_PEAKS.json (+ a TF32 cuBLAS
              peak measured in this run for the tf32 engines)
  cpu_baseline : the oracle's FAISS-CPU-like fp32 search ("port") on this box's host cores, bounded sample
  parity    : rank 0 checks the first queries of the timed result against the CPU oracle (all N)
  sweep     : other batch sizes; `extras`: L2 metric, the 3xTF32 engine, clustered / duplicate-laden corpora
  scale_configs : BASELINE.json configs[0], [4], [2] and [3] at FULL size (every N; --scale-configs picks), each with parity

`--impl reference` times the CPU path alone (rank 0 only) and prints the same line with "impl": "reference".
"""
import argparse
import csv
import glob
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
FALLBACK_PEAKS = {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}
BLK = 65536            # rows are generated per block of this many rows: the same values whatever the sharding
N_CENTRES = 1024       # --dist mixture / dup
SIGMA = 0.3            # norm of the perturbation around a unit-norm centre (0.3 / sqrt(d) per coordinate)


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return d, "measured"
    return dict(FALLBACK_PEAKS), "fallback"


def ncu_traffic(mode, batch, rows, world):
    """dram__bytes_read.sum + dram__bytes_write.sum of the main scoring launch, read from the committed
    `ncu --set full` capture of this configuration (profiles/rNN_ncu_full_<mode>_b<batch>_main.csv, newest
    round first); None when there is no capture for it."""
    if rows != N_ROWS or world != 1:
        return None, None
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    for path in sorted(glob.glob(os.path.join(ROOT, "profiles", f"r*_ncu_full_{mode}_b{batch}_main.csv")), reverse=True):
        try:
            with open(path, newline="") as f:
                rows_csv = list(csv.reader(f))
            hdr, units, vals = rows_csv[0], rows_csv[1], rows_csv[2]
            total = 0.0
            for name in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                i = hdr.index(name)
                total += float(vals[i]) * scale[units[i]]
            return total, os.path.relpath(path, ROOT)
        except Exception:
            continue
    return None, None


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


# ----------------------------------------------------------------------------------------------------
# synthetic data (torch Philox streams keyed by (seed, block): identical rows on every rank and in both arms)
# ----------------------------------------------------------------------------------------------------
def _block(torch, seed, b, d, device, dist):
    """Block b (rows [b*BLK, (b+1)*BLK)) of the stream `seed`, un-normalised, fp32 on `device`."""
    g = torch.Generator(device=device)
    g.manual_seed(seed * 1_000_003 + b)
    rows = torch.randn((BLK, d), generator=g, dtype=torch.float32, device=device)
    if dist == "gauss":
        return rows
    gc = torch.Generator(device=device)
    gc.manual_seed(7_777_777 + d)
    centres = torch.randn((N_CENTRES, d), generator=gc, dtype=torch.float32, device=device)
    centres /= centres.norm(dim=1, keepdim=True)
    which = torch.randint(0, N_CENTRES, (BLK,), generator=g, device=device)
    rows = centres[which] + rows * (SIGMA / d ** 0.5)
    if dist == "dup":        # 1 % of the rows are exact copies of the row 50 places earlier (same block)
        i = torch.arange(99, BLK, 100, device=device)
        rows[i] = rows[i - 50]
    return rows


def gen_rows(torch, seed, row0, nrows, d, device, dist="gauss"):
    """Rows [row0, row0 + nrows) of the stream `seed` (un-normalised)."""
    out = torch.empty((nrows, d), dtype=torch.float32, device=device)
    b0 = row0 // BLK * BLK
    while b0 < row0 + nrows:
        rows = _block(torch, seed, b0 // BLK, d, device, dist)
        lo, hi = max(b0, row0), min(b0 + BLK, row0 + nrows)
        out[lo - row0:hi - row0] = rows[lo - b0:hi - b0]
        b0 += BLK
    return out


def gen_queries(torch, b, d, device, dist, nrows):
    """Queries: i.i.d. Gaussian for --dist gauss; perturbed corpus rows (same sigma) for mixture / dup."""
    if dist == "gauss":
        return gen_rows(torch, 1, 0, b, d, device, "gauss")
    g = torch.Generator(device=device)
    g.manual_seed(424_242)
    src = torch.randint(0, nrows, (b,), generator=g, device=device)
    q = torch.empty((b, d), dtype=torch.float32, device=device)
    src_h = src.cpu().tolist()
    cache = {}
    for i, r in enumerate(src_h):                       # at most b blocks are generated
        blk = r // BLK
        if blk not in cache:
            if len(cache) > 4:
                cache.clear()
            rows = _block(torch, 0, blk, d, device, dist)
            cache[blk] = rows / rows.norm(dim=1, keepdim=True)
        q[i] = cache[blk][r % BLK]
    q += torch.randn((b, d), generator=g, dtype=torch.float32, device=device) * (SIGMA / d ** 0.5)
    return q


# ----------------------------------------------------------------------------------------------------
# CPU port (FAISS-flat restatement), threads chosen by a short calibration
# ----------------------------------------------------------------------------------------------------
def cpu_search(fo, x_host, q_host, k, metric, threads):
    from threadpoolctl import threadpool_limits
    with threadpool_limits(limits=threads):
        t0 = time.perf_counter()
        fo.search_fp32_blas(x_host, q_host, k, metric, bs_q=4096, bs_x=16384)
        return time.perf_counter() - t0


def cpu_pick_threads(fo, x_host, q_host, k, metric, cores):
    """Best of {1, cores/2, cores} BLAS/OpenMP threads on a short slice (thread thrash is real: under torchrun
    OMP_NUM_THREADS=1 is exported and one thread once beat 32)."""
    best, best_t = None, None
    xs, qs = x_host[:131072], q_host[:min(len(q_host), 256)]
    for t in sorted({1, max(1, cores // 2), cores}):
        cpu_search(fo, xs[:16384], qs, k, metric, t)      # spin the pool up
        sec = cpu_search(fo, xs, qs, k, metric, t)
        if best_t is None or sec < best_t:
            best, best_t = t, sec
    return best


def cpu_baseline_run(fo, x_host, q_host, k, metric, budget_s=20.0):
    cores = len(os.sched_getaffinity(0))
    threads = cpu_pick_threads(fo, x_host, q_host, k, metric, cores)
    # bounded sample: as many of the batch's queries as ~budget_s allows, over the FULL corpus, one pass
    probe_q = min(len(q_host), 128)
    sec_probe = cpu_search(fo, x_host, q_host[:probe_q], k, metric, threads)
    nqs = int(min(len(q_host), max(probe_q, probe_q * budget_s / max(sec_probe, 1e-3))))
    nqs = max(probe_q, nqs // 128 * 128) if nqs < len(q_host) else len(q_host)
    sec = cpu_search(fo, x_host, q_host[:nqs], k, metric, threads) if nqs != probe_q else sec_probe
    return {"value": nqs / sec, "unit": "queries/s", "cores": threads, "kind": "port",
            "host_cores_available": cores, "blas_threads": threads,
            "OPENBLAS_NUM_THREADS": os.environ.get("OPENBLAS_NUM_THREADS"), "OMP_NUM_THREADS": os.environ.get("OMP_NUM_THREADS"),
            "sample": f"{nqs} of the {len(q_host)} queries, one pass over the full {x_host.shape[0]}x{x_host.shape[1]} corpus "
                      f"({sec:.2f} s, sgemm blocks {min(nqs, 4096)}x16384, best of 1/{max(1, cores // 2)}/{cores} threads); "
                      "CPU restatement of FAISS-flat, not FAISS"}, sec, nqs


def host_corpus(torch, b2r, rows, d, dist, device):
    """The full corpus in host memory, identical to what the GPU arm indexes (generated on the device in blocks,
    normalised by kernel (a), copied back); NumPy fallback without a device (reference arm only)."""
    x_host = np.empty((rows, d), dtype=np.float32)
    if device is not None:
        chunk = 131072
        for r0 in range(0, rows, chunk):
            blk = gen_rows(torch, 0, r0, min(chunk, rows - r0), d, device, dist)
            b2r.normalize_L2(blk)
            x_host[r0:r0 + blk.shape[0]] = blk.cpu().numpy()
        return x_host, "torch Philox streams on the device + kernel (a) normalisation (same rows as the GPU arm)"
    from oracle import flat_oracle as fo
    for r0 in range(0, rows, BLK):
        x_host[r0:r0 + BLK] = fo.synth_rows(0, r0, min(BLK, rows - r0), d)
    x_host /= np.linalg.norm(x_host, axis=1, keepdims=True)
    return x_host, "NumPy Philox streams (no CUDA device visible to the reference arm: not the GPU arm's rows)"


def run_reference(args, rank, world):
    """The reference arm: the CPU restatement of FAISS-flat search on this box's host cores, over the same rows and
    queries as the GPU arm (the device is used to GENERATE them only; no kernel of the engine is on the timed path)."""
    if rank != 0:
        return
    import torch
    from oracle import flat_oracle as fo
    fo.build()
    device, b2r = None, None
    if torch.cuda.is_available():
        try:
            import rag_tl_domainllm_optimizer_b200 as b2r
            device = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
        except Exception:
            device = None
    x, how = host_corpus(torch, b2r, args.rows, DIM, args.dist, device)
    if device is not None:
        qd = gen_queries(torch, args.batch, DIM, device, args.dist, args.rows)
        b2r.normalize_L2(qd)
        q = qd.cpu().numpy()
        del qd
        torch.cuda.empty_cache()
    else:
        q = fo.synth_rows(1, 0, args.batch, DIM)
        q /= np.linalg.norm(q, axis=1, keepdims=True)
    metric = 1 if args.metric == "l2" else 0
    cores = len(os.sched_getaffinity(0))
    threads = cpu_pick_threads(fo, x, q, args.k, metric, cores)
    nqs = min(args.batch, args.cpu_queries)
    for _ in range(min(args.warmup, 1)):
        cpu_search(fo, x, q[:nqs], args.k, metric, threads)
    t = [cpu_search(fo, x, q[:nqs], args.k, metric, threads) for _ in range(args.steps)]
    sec = sum(t) / len(t)
    qps = nqs / sec
    sample = (f"{nqs} of the {args.batch} queries per step over the full {args.rows}x{DIM} corpus, sgemm blocks "
              f"{nqs}x16384, {threads} BLAS/OpenMP threads (best of 1/{max(1, cores // 2)}/{cores})")
    line = {
        "impl": "reference", "metric": "flat top-k search queries/sec", "value": qps, "unit": "queries/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec * 1e3,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, world),
        "cpu_baseline": {"value": qps, "unit": "queries/s", "cores": threads, "kind": "port", "sample": sample,
                         "host_cores_available": cores, "OPENBLAS_NUM_THREADS": os.environ.get("OPENBLAS_NUM_THREADS"),
                         "OMP_NUM_THREADS": os.environ.get("OMP_NUM_THREADS"), "rows_from": how},
        "e2e": {"value": qps, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "CPU restatement of FAISS-flat (NumPy/OpenBLAS sgemm + C heap), not FAISS: the reference pins no "
                "FAISS version and FAISS is not installable here (BASELINE.md §3)",
    }
    print(json.dumps(line), flush=True)


def workload_config(args, world):
    return {"workload": f"{'configs[1]: ' if (args.rows, DIM) == (N_ROWS, 768) else ''}{args.rows}x{DIM} corpus, query batch {args.batch}, k={args.k}, "
                        f"{'exact fp32' if args.mode == 'exact' else 'bf16'} {'squared-L2' if args.metric == 'l2' else 'inner-product'} flat search",
            "rows": args.rows, "dim": DIM, "batch": args.batch, "k": args.k, "metric": args.metric, "dist": args.dist,
            "mode": args.mode + ("+bf16-shadow prefilter" if getattr(args, "shadow", False) and args.mode == "exact" else ""),
            "sharding": f"rows/{world}" if world > 1 else "none",
            "l2": "inputs larger than L2: corpus shard re-read from HBM every step (>= 126 MB)"}


def measure_tf32_peak(torch, device, seconds=2.0):
    """cuBLAS TF32 8192^3 (fp32 tensors, allow_tf32): best of 10 (burst) and back to back (sustained) — the recipe of
    MEASURED_PEAKS.json's bf16 figure, for the kind::tf32 engines that file has no number for."""
    old = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = True
    try:
        n = 8192
        a = torch.randn((n, n), device=device)
        b = torch.randn((n, n), device=device)
        for _ in range(3):
            a @ b
        torch.cuda.synchronize()
        best = 1e9
        for _ in range(10):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            a @ b
            e1.record()
            e1.synchronize()
            best = min(best, e0.elapsed_time(e1))
        reps = max(10, int(seconds * 1e3 / best))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            a @ b
        e1.record()
        e1.synchronize()
        flop = 2.0 * n ** 3
        return {"tf32_tflops": flop / (best * 1e-3) / 1e12, "tf32_tflops_sustained": flop * reps / (e0.elapsed_time(e1) * 1e-3) / 1e12,
                "how": f"torch.matmul fp32 {n}^3 with allow_tf32 (cuBLAS TF32): best of 10, and {reps} back to back"}
    finally:
        torch.backends.cuda.matmul.allow_tf32 = old


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
    ap.add_argument("--metric", default="ip", choices=["ip", "l2"])
    ap.add_argument("--dist", default="gauss", choices=["gauss", "mixture", "dup"],
                    help="corpus: i.i.d. Gaussian | 1024-centre mixture, queries = perturbed rows | mixture + 1%% duplicate rows")
    ap.add_argument("--engine", type=int, default=0)
    ap.add_argument("--shadow", action="store_true", help="exact mode: score a bf16 shadow copy, re-score fp32 rows")
    ap.add_argument("--sweep", default="1,8,32,128,1024", help="comma list of extra batch sizes ('' = none)")
    ap.add_argument("--extras", default="l2,engine3,mixture,dup,bf16", help="extra batch-1024-style rows ('' = none)")
    ap.add_argument("--scale-configs", default=None, help="comma list of c5,c3 ('' = none; default: both when --gpus > 1)")
    ap.add_argument("--cpu-queries", type=int, default=256)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-peaks", action="store_true")
    ap.add_argument("--check-queries", type=int, default=32)
    args = ap.parse_args()
    DIM = args.dim
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else max(args.warmup, 1)

    T_START = time.perf_counter()
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

    def sum_over_ranks(v: float) -> float:
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    peaks, peak_src = load_peaks()
    chunk = 131072

    # ---- one configuration: build the (sharded) index ------------------------------------------
    def build_index(rows, d, mode, metric, dist_name, engine=0, shadow=False):
        storage = "f32" if mode == "exact" else "bf16"
        lo, hi = b2r.shard_bounds(rows, world, rank)
        sidx = b2r.ShardedIndexFlat(d, b2r.METRIC_L2 if metric == "l2" else b2r.METRIC_INNER_PRODUCT, storage=storage,
                                    device=local_rank, shadow=shadow and storage == "f32")
        sidx.local.set_engine(engine)
        sidx.local.reserve(hi - lo)
        ing0, ing1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ingest_ms = 0.0
        for r0 in range(lo, hi, chunk):
            blk = gen_rows(torch, 0, r0, min(chunk, hi - r0), d, device, dist_name)
            b2r.normalize_L2(blk)                      # kernel (a), in place, as faiss.normalize_L2 is used
            ing0.record()
            sidx.add_local(blk, lo, rows)              # kernel (a) again: cast / copy into the index, row norms
            ing1.record()
            ing1.synchronize()
            ingest_ms += ing0.elapsed_time(ing1)
            del blk
        torch.cuda.synchronize()
        return sidx, lo, hi, ingest_ms

    def make_queries(b, d, dist_name, rows):
        q = gen_queries(torch, b, d, device, dist_name, rows)
        b2r.normalize_L2(q)
        return q

    def timed_device(sidx, q, k, steps, warmup):
        for _ in range(warmup):
            sidx.search(q, k)
        sidx.local.score_time(reset=True)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            D, I = sidx.search(q, k)
        e1.record()
        barrier()
        ms = max_over_ranks(e0.elapsed_time(e1))
        tot, cnt = sidx.local.score_time(reset=True)
        return ms / steps, (tot / cnt if cnt else 0.0), D, I

    def timed_e2e(sidx, q_host, k, steps, warmup):
        for _ in range(warmup):
            sidx.search(q_host, k)
        barrier()
        t0 = time.perf_counter()
        for _ in range(steps):
            D, I = sidx.search(q_host, k)              # host buffers: H2D of the queries, D2H of D/I inside the call
            D_h = D.cpu() if hasattr(D, "cpu") else torch.from_numpy(D)
            I_h = I.cpu() if hasattr(I, "cpu") else torch.from_numpy(I)
        barrier()
        return max_over_ranks((time.perf_counter() - t0) * 1e3) / steps, D_h, I_h

    def stats_all_ranks(sidx):
        st = sidx.local.last_stats()
        out = dict(st)
        for key in ("fallback_queries", "exhaustive_queries", "unverified"):
            out[key] = int(sum_over_ranks(float(st[key])))
        return out

    # ---- parity: candidates by an fp64 product, scores and order by the oracle's defined arithmetic ----------
    def parity_check(sidx, lo, hi, rows, d, mode, metric, dist_name, q_dev, k, D_got, I_got, nc, host_rows_ok):
        """Rank 0 compares the engine's answer for the first `nc` queries with the oracle.
        Small corpora (host_rows_ok): the oracle runs entirely on the host over rows regenerated there
        (oracle.search_ref: chunked fp64 BLAS candidates + DOT64 re-scoring in C).
        Full-size configs: the fp64 candidate product runs on each rank's GPU over ITS shard (torch fp64 matmul
        + topk, k + 64 candidates), the candidates' rows travel to rank 0 and the oracle's C code re-scores and
        orders them — ids AND score bits must match."""
        from oracle import flat_oracle as fo
        metric_i = 1 if metric == "l2" else 0
        nc = min(nc, q_dev.shape[0])
        qh = q_dev[:nc].cpu().numpy()
        Dg = D_got[:nc].cpu().numpy() if hasattr(D_got, "cpu") else np.asarray(D_got)[:nc]
        Ig = I_got[:nc].cpu().numpy() if hasattr(I_got, "cpu") else np.asarray(I_got)[:nc]
        if host_rows_ok:
            if rank != 0:
                return None
            fo.build()
            xh, _ = host_corpus(torch, b2r, rows, d, dist_name, device)
            xo, qo = (fo.round_bf16(xh), fo.round_bf16(qh)) if mode == "bf16" else (xh, qh)
            D0, I0 = fo.search_ref(xo, qo, k, metric_i)
            method = "oracle.search_ref on the host over the regenerated full corpus"
            res = {"recall_at_k_vs_fp32": None}
            if mode == "bf16":
                _, If = fo.search_ref(xh, qh, k, metric_i)
                res["recall_at_k_vs_fp32"] = float(np.mean([len(set(a) & set(b)) / k for a, b in zip(Ig, If)]))
            del xh
        else:
            extra = 64
            kc = min(hi - lo, k + extra)
            q64 = q_dev[:nc].double()
            if mode == "bf16":
                q64 = q_dev[:nc].bfloat16().double()
            best_s = torch.full((nc, kc), -float("inf"), dtype=torch.float64, device=device)
            best_i = torch.zeros((nc, kc), dtype=torch.int64, device=device)
            step_rows = max(16384, (1 << 28) // d)      # fp64 copy of a chunk stays near 2 GB
            for r0 in range(lo, hi, step_rows):
                m = min(step_rows, hi - r0)
                xb = torch.empty((m, d), dtype=torch.float32, device=device)
                sidx.local_export_rows(r0 - lo, m, xb)                      # the STORED rows (bf16 mode: rounded values)
                x64 = xb.double()
                s = q64 @ x64.T
                if metric_i == 1:
                    s = 2.0 * s - (x64 * x64).sum(1)[None, :]
                ids = torch.arange(r0, r0 + m, device=device)[None, :].expand(nc, m)
                s = torch.cat([best_s, s], 1)
                ids = torch.cat([best_i, ids], 1)
                top = torch.topk(s, kc, dim=1)
                best_s, best_i = top.values, torch.gather(ids, 1, top.indices)
                del xb, x64, s, ids
            # candidate rows of this shard, then everything to every rank (tiny) and the oracle on rank 0
            rows_c = torch.empty((nc * kc, d), dtype=torch.float32, device=device)
            sidx.local_gather_rows((best_i.flatten() - lo).clamp_(0, hi - lo - 1), rows_c)
            if world > 1:
                all_s = [torch.empty_like(best_s) for _ in range(world)]
                all_i = [torch.empty_like(best_i) for _ in range(world)]
                all_r = [torch.empty_like(rows_c) for _ in range(world)]
                dist.all_gather(all_s, best_s)
                dist.all_gather(all_i, best_i)
                dist.all_gather(all_r, rows_c)
                best_s, best_i, rows_c = torch.cat(all_s, 1), torch.cat(all_i, 1), torch.cat([r.view(nc, kc, d) for r in all_r], 1)
            else:
                rows_c = rows_c.view(nc, kc, d)
            if rank != 0:
                return None
            fo.build()
            kk = best_s.shape[1]
            order = torch.argsort(best_s, dim=1, descending=True)[:, :min(kk, k + extra)]
            ids_h = torch.gather(best_i, 1, order).cpu().numpy()
            rows_h = torch.gather(rows_c, 1, order[:, :, None].expand(-1, -1, d)).cpu().numpy()
            qo = fo.round_bf16(qh) if mode == "bf16" else qh
            D0 = np.empty((nc, k), dtype=np.float32)
            I0 = np.empty((nc, k), dtype=np.int64)
            for i in range(nc):                        # oracle re-scoring over the candidate rows of query i
                D0[i], I0[i] = fo.rescore_rows(rows_h[i], ids_h[i], qo[i], k, metric_i)
            method = (f"fp64 candidates (k+{extra} per shard) by torch matmul/topk on each rank's stored rows; "
                      "scores and order by oracle DOT64 re-scoring (C) on rank 0")
            res = {"recall_at_k_vs_fp32": None}
        return {"checked_queries": int(nc), "ids_identical": bool(np.array_equal(Ig, I0)),
                "scores_bit_identical": bool(np.array_equal(Dg.view(np.uint32), D0.view(np.uint32))),
                "method": method, **res}

    # ---- one measured configuration -------------------------------------------------------------
    def run_config(rows, d, batch, k, mode, metric, dist_name, engine, steps, warmup, *, shadow=False, e2e=True,
                   parity_queries=0, host_rows_ok=True, keep=False, sample_clocks=False):
        sidx, lo, hi, ingest_ms = build_index(rows, d, mode, metric, dist_name, engine, shadow)
        storage = "f32" if mode == "exact" else "bf16"
        esize = 4 if storage == "f32" else 2
        q = make_queries(batch, d, dist_name, rows)
        sampler = ClockSampler(local_rank) if sample_clocks else None
        if sampler:
            sampler.start()
        ms_step, score_ms, D, I = timed_device(sidx, q, k, steps, warmup)
        clocks = sampler.stop() if sampler else None
        st = stats_all_ranks(sidx)
        res = {"rows": rows, "dim": d, "batch": batch, "k": k, "mode": mode, "metric": metric, "dist": dist_name,
               "engine": {1: "fma", 2: "umma", 3: "umma-3xtf32", 4: "umma-bf16-shadow"}.get(st["engine"], st["engine"]),
               "qps": batch / (ms_step * 1e-3), "ms_per_step": ms_step, "score_ms": score_ms,
               "fallback_queries": st["fallback_queries"], "exhaustive_queries": st["exhaustive_queries"],
               "unverified_queries": st["unverified"], "kprime": st["kprime"], "launches": st["launches"]}
        n_local = hi - lo
        rows_k = (st["main_rows"] or n_local) if st["engine"] in (2, 3, 4) else n_local   # rows the timed launch covers
        esize_score = 2 if st["engine"] == 4 else esize
        res["_roof"] = (rows_k, esize_score, st["engine"], n_local)
        if score_ms:
            res["hbm_gbs"] = rows_k * d * esize_score / (score_ms * 1e-3) / 1e9
            res["tflops"] = 2.0 * batch * rows_k * d / (score_ms * 1e-3) / 1e12
            res["main_launch_share"] = score_ms / ms_step
        if clocks:
            res["clocks"] = clocks
        if ingest_ms:
            res["ingest_gbs"] = n_local * d * (4 + esize) / (ingest_ms * 1e-3) / 1e9
        if e2e:
            q_pinned = torch.empty((batch, d), dtype=torch.float32, pin_memory=True)
            q_pinned.copy_(q)
            e2e_ms, _, _ = timed_e2e(sidx, q_pinned.numpy(), k, steps, warmup)
            res["e2e_qps"], res["e2e_ms"] = batch / (e2e_ms * 1e-3), e2e_ms
        if parity_queries:
            res["parity"] = parity_check(sidx, lo, hi, rows, d, mode, metric, dist_name, q, k, D, I, parity_queries, host_rows_ok)
            if res["parity"] is not None:
                res["parity"]["unverified_queries"] = st["unverified"]
        if keep:
            return res, sidx, q, (lo, hi)
        del sidx, q
        torch.cuda.empty_cache()
        return res

    # helpers the parity check needs from the index (stored rows by range / by id), through the C ABI
    def _export_rows(self, row0, n, out):
        from rag_tl_domainllm_optimizer_b200 import _lib
        import ctypes
        buf = out if self.local.storage == "f32" else torch.empty(out.shape, dtype=torch.bfloat16, device=out.device)
        _lib.check(_lib.lib().b2r_index_export_rows(self.local._h, int(row0), int(n), ctypes.c_void_p(buf.data_ptr()),
                                                    torch.cuda.current_stream().cuda_stream))
        if buf is not out:
            out.copy_(buf.float())

    def _gather_rows(self, ids, out):
        ids_h = ids.cpu().tolist()
        one = torch.empty((1, out.shape[1]), dtype=torch.float32, device=out.device)
        for j, r in enumerate(ids_h):
            _export_rows(self, r, 1, one)
            out[j] = one[0]

    b2r.ShardedIndexFlat.local_export_rows = _export_rows
    b2r.ShardedIndexFlat.local_gather_rows = _gather_rows

    # ---- headline ----------------------------------------------------------------------------
    head, sidx, q, (lo, hi) = run_config(args.rows, DIM, args.batch, args.k, args.mode, args.metric, args.dist, args.engine,
                                         args.steps, args.warmup, shadow=args.shadow, parity_queries=0, keep=True,
                                         sample_clocks=True)
    qps, ms_step, score_ms = head["qps"], head["ms_per_step"], head["score_ms"]
    rows_k, esize_score, eng, n_local = head["_roof"]
    storage = "f32" if args.mode == "exact" else "bf16"

    # ---- roofline of the dominant (scoring) kernel -------------------------------------------
    tf32 = None
    if not args.no_peaks and eng in (2, 3):
        tf32 = measure_tf32_peak(torch, device)
    bytes_alg = rows_k * DIM * esize_score
    flops_alg = 2.0 * args.batch * rows_k * DIM
    hbm_gbs = bytes_alg / (score_ms * 1e-3) / 1e9 if score_ms else 0.0
    tflops = flops_alg / (score_ms * 1e-3) / 1e12 if score_ms else 0.0
    hbm_frac = hbm_gbs / peaks["hbm_gbs"]
    # the timed region is tens of ms (no power-cap settling): compare with the BURST cuBLAS figure, the harder one;
    # the fraction of the sustained figure is reported beside it.  bf16 MMAs (fast mode, shadow) against the bf16
    # peak of MEASURED_PEAKS.json; kind::tf32 MMAs (exact mode) against the TF32 cuBLAS peak measured in this run
    bf16_peak = peaks["bf16_tflops"]
    bf16_sustained = peaks.get("bf16_tflops_sustained", bf16_peak)
    if eng in (2, 3) and storage == "f32" and tf32:
        tc_peak, tc_sustained, tc_name = tf32["tf32_tflops"], tf32["tf32_tflops_sustained"], "tf32 cuBLAS, measured in this run"
    else:
        tc_peak, tc_sustained, tc_name = bf16_peak, bf16_sustained, f"bf16 cuBLAS, {peak_src}"
    tc_frac = tflops / tc_peak
    if hbm_frac >= tflops / bf16_peak and hbm_frac >= tc_frac:
        roof = {"bound": "hbm", "achieved": hbm_gbs, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": hbm_frac}
    else:
        roof = {"bound": "tensor", "achieved": tflops, "peak": tc_peak, "unit": "TFLOP/s", "frac": tc_frac,
                "peak_kind": tc_name, "frac_of_sustained_peak": tflops / tc_sustained,
                "frac_of_bf16_burst_peak": tflops / bf16_peak}
    if tf32:
        roof["tf32_peak_measured"] = tf32
    traffic, traffic_src = ncu_traffic(args.mode, args.batch, args.rows, world)
    kernel_name = {2: "score_umma_kernel<tf32> (main launch)", 3: "score_umma_kernel<3xtf32> (main launch)",
                   4: "score_umma_kernel<bf16 shadow> (main launch)"}.get(eng, "score_fma_kernel (all passes)")
    if storage == "bf16" and eng == 2:
        kernel_name = "score_umma_kernel<bf16> (main launch)"
    roof.update({"traffic": traffic, "traffic_source": traffic_src or "no committed ncu --set full capture for this configuration",
                 "kernel": kernel_name, "rows_per_launch": rows_k, "kernel_ms": score_ms,
                 "peak_source": f"{peak_src} MEASURED_PEAKS.json (hbm copy {peaks['hbm_gbs']} GB/s; bf16 burst {bf16_peak} / sustained {bf16_sustained} TFLOP/s)",
                 "algorithmic": {"bytes_per_launch_set": bytes_alg, "flops_per_launch_set": flops_alg}})

    # ---- e2e already measured; parity spot-check (all N) and CPU baseline (N = 1) on rank 0 -----------------
    D_chk, I_chk = sidx.search(q, args.k)
    st_head = stats_all_ranks(sidx)
    parity = None
    if not args.no_cpu:
        parity = parity_check(sidx, lo, hi, args.rows, DIM, args.mode, args.metric, args.dist, q, args.k, D_chk, I_chk,
                              args.check_queries, host_rows_ok=args.rows <= 2_000_000)
        if parity is not None:
            parity["unverified_queries"] = st_head["unverified"]
            parity["fallback_queries"] = st_head["fallback_queries"]
    cpu_base = None
    if rank == 0 and world == 1 and not args.no_cpu:
        from oracle import flat_oracle as fo
        fo.build()
        xh, _ = host_corpus(torch, b2r, args.rows, DIM, args.dist, device)
        cpu_base, _, _ = cpu_baseline_run(fo, xh, q.cpu().numpy(), args.k, 1 if args.metric == "l2" else 0)
        del xh
    del sidx
    torch.cuda.empty_cache()

    # ---- sweep over the other batch sizes (1 GPU only) ------------------------------------------
    sweep = []
    if world == 1 and args.sweep:
        sidx2, _, _, _ = build_index(args.rows, DIM, args.mode, args.metric, args.dist, args.engine, args.shadow)
        for b in [int(v) for v in args.sweep.split(",") if v]:
            if b == args.batch:
                sweep.append({"batch": b, "mode": args.mode, "qps": qps, "score_ms": score_ms,
                              "hbm_frac": hbm_frac, "tensor_frac": tc_frac})
                continue
            qb = make_queries(b, DIM, args.dist, args.rows)
            steps = max(5, min(args.steps, 50))
            ms_b, sc_b, _, _ = timed_device(sidx2, qb, args.k, steps, 3)
            sweep.append({"batch": b, "mode": args.mode, "qps": b / (ms_b * 1e-3), "score_ms": sc_b,
                          "hbm_frac": (bytes_alg / (sc_b * 1e-3) / 1e9) / peaks["hbm_gbs"] if sc_b else 0.0,
                          "tensor_frac": (2.0 * b * rows_k * DIM / (sc_b * 1e-3) / 1e12) / tc_peak if sc_b else 0.0})
        del sidx2
        torch.cuda.empty_cache()

    # ---- extra rows at the headline shape: other metric / engine / distributions / mode (1 GPU only) ---------
    extras = []
    if world == 1 and args.extras:
        steps_x = max(5, min(args.steps, 10))
        for name in [v for v in args.extras.split(",") if v]:
            cfg = {"mode": args.mode, "metric": args.metric, "dist": args.dist, "engine": args.engine}
            if name == "l2":
                cfg["metric"] = "l2"
            elif name == "engine3":
                if args.mode != "exact":
                    continue
                cfg["engine"] = 3
            elif name in ("mixture", "dup"):
                cfg["dist"] = name
            elif name == "bf16":
                cfg["mode"] = "bf16"
                cfg["engine"] = 0
            else:
                continue
            if all(cfg[key] == getattr(args, key) for key in cfg):
                continue
            r = run_config(args.rows, DIM, args.batch, args.k, cfg["mode"], cfg["metric"], cfg["dist"], cfg["engine"], steps_x, 3,
                           e2e=False, parity_queries=8 if not args.no_cpu else 0, host_rows_ok=args.rows <= 2_000_000)
            r.pop("_roof", None)
            r["name"] = name
            r["tensor_frac_bf16_burst"] = r.get("tflops", 0.0) / bf16_peak
            extras.append(r)

    # ---- the configs north_star puts its scaling target on, at full size --------------------------------
    scale = []
    # default: both, at every N (the driver's 1/2/4/8 scaling run then carries them), unless the run is already long
    t_run = max_over_ranks(time.perf_counter() - T_START)      # the same decision on every rank
    which = args.scale_configs if args.scale_configs is not None else ("c1,c5,c3,c4" if t_run < 400 else "")
    for name in [v for v in which.split(",") if v]:
        if name == "c5":      # configs[4]: 100M x 768 bf16, batch 4096, k=10
            shape = dict(rows=100_000_000, d=768, batch=4096, k=10, mode="bf16")
        elif name == "c3":    # configs[2]: 10M x 1024, batch 256, k=100
            shape = dict(rows=10_000_000, d=1024, batch=256, k=100, mode="exact")
        elif name == "c4":    # configs[3]: 5M x 4096, 256 queries, k=4
            shape = dict(rows=5_000_000, d=4096, batch=256, k=4, mode="bf16")
        elif name == "c1":    # configs[0]: 10k x 384, 64 queries, k=5, exact (launch-bound: q/s and us per search)
            shape = dict(rows=10_000, d=384, batch=64, k=5, mode="exact")
        else:
            continue
        try:
            r = run_config(shape["rows"], shape["d"], shape["batch"], shape["k"], shape["mode"], "ip", "gauss", 0,
                           steps=200 if name == "c1" else 5, warmup=3, e2e=False, parity_queries=8 if not args.no_cpu else 0,
                           host_rows_ok=shape["rows"] <= 2_000_000, sample_clocks=True)
            r["us_per_search"] = r["ms_per_step"] * 1e3
            r.pop("_roof", None)
            r["tensor_frac_bf16_burst"] = r.get("tflops", 0.0) / bf16_peak
            r["hbm_frac"] = r.get("hbm_gbs", 0.0) / peaks["hbm_gbs"]
        except Exception as exc:   # e.g. a corpus that does not fit this many GPUs: say so instead of shrinking it
            r = {"error": f"{type(exc).__name__}: {exc}"[:300]}
            torch.cuda.empty_cache()
        r["config"] = name
        scale.append(r)

    if rank == 0:
        line = {
            "metric": "flat top-k search queries/sec", "value": qps, "unit": "queries/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None,
            "dtype": {2: "tf32", 3: "tf32", 4: "bf16"}.get(eng, "f32") if storage == "f32" else "bf16",
            "data": "synthetic", "config": workload_config(args, world), "clocks": head.get("clocks"),
            "e2e": {"value": head["e2e_qps"], "unit": "queries/s", "ms_per_step": head["e2e_ms"],
                    "h2d_bytes_per_step": args.batch * DIM * 4, "d2h_bytes_per_step": args.batch * args.k * 12},
            "gpu_launches": head["launches"] * args.steps,
            "roofline": roof, "cpu_baseline": cpu_base, "parity": parity, "sweep": sweep, "extras": extras,
            "scale_configs": scale,
            "engine": head["engine"], "kprime": head["kprime"],
            "fallback_queries": head["fallback_queries"], "exhaustive_queries": head["exhaustive_queries"],
            "ingest": {"kernel": "ingest_rows_kernel", "gbs": head.get("ingest_gbs", 0.0),
                       "frac_of_hbm_peak": head.get("ingest_gbs", 0.0) / peaks["hbm_gbs"],
                       "bytes_per_row": DIM * (4 + (4 if storage == "f32" else 2))},
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
