// score_exact.cuh — the scan of last resort: exhaustive DOT64 scoring of the whole shard for the few queries
// that neither the first pass's proof nor the range pass could settle (thousands of rows tied with the k-th
// score: exact duplicates, constant corpora).  It computes, for every stored row, the very fp64 reduction that
// defines a returned score (common.cuh: warp_reduce64 order), so there is no error bound and nothing to prove:
// the top-k it returns is the definition.  What a CPU faiss IndexFlat does for every query, done here for the
// pathological ones only.
//
// One cooperative launch (all CTAs resident, software grid barrier), 8 warps per CTA:
//   phase A  for each group of 8 listed queries: warp = one row per iteration (consecutive warps take consecutive
//            rows), lane = its DOT64 lane, 8 fp64 accumulators (one per query), 8 butterflies; lane j then owns
//            query j's exact key; keys that beat the CTA's running k-th best are appended to a shared-memory
//            buffer, compacted by a block bitonic sort when nearly full; each CTA leaves its best k per query in
//            global memory;
//   barrier
//   phase B  the wave's queries are dealt to the CTAs; each merges the per-CTA lists (gather_topk) and writes
//            D/I (or packed keys) over the earlier passes' rows;
//   barrier, next wave (the per-CTA lists are sized for `wave` queries).
// The launch returns at once when the list is empty, which is the normal case.
// HBM bytes: rows * ld * sizeof(elem) per group of 8 queries; fp64 FMAs: 8 * rows * d per group.
#pragma once
#include "common.cuh"
#include "finalize.cuh"

namespace b2r {

constexpr int kExQB = 8;          // queries per scan
constexpr int kExThreads = 256;   // 8 warps: 8 rows per CTA iteration

__device__ __forceinline__ void grid_barrier(uint32_t* bar, uint32_t& epoch) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    epoch += gridDim.x;
    atomicAdd(bar, 1u);
    uint32_t seen;
    do {
      asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(seen) : "l"(bar) : "memory");
      if (seen < epoch) __nanosleep(128);
    } while (seen < epoch);
    __threadfence();
  }
  __syncthreads();
}

template <typename T>
__global__ void __launch_bounds__(kExThreads) score_exact_kernel(FinalizeParams p, uint64_t* __restrict__ lists,
                                                                  int wave, int cap) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  uint64_t* buf = reinterpret_cast<uint64_t*>(smem_raw);   // phase A: [kExQB][cap]; phase B: [sortn] + [next_pow2(k)]
  __shared__ int cnt_s[kExQB];
  __shared__ unsigned long long thr_s[kExQB];
  __shared__ int fill_s;
  __shared__ int off_s[kFinMaxSlots + 1];
  const int n2 = int(min(*p.ex_n, uint32_t(p.nq)));
  if (n2 == 0) return;                                      // nothing left unproven: the normal case
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int grid = gridDim.x;
  const T* X = reinterpret_cast<const T*>(p.X);
  uint32_t epoch = 0;
  const int64_t iters = (p.n + int64_t(grid) * 8 - 1) / (int64_t(grid) * 8);

  for (int w0 = 0; w0 < n2; w0 += wave) {
    const int nw = min(wave, n2 - w0);
    // ---------------- phase A: scan ----------------
    for (int g0 = 0; g0 < nw; g0 += kExQB) {
      const int nqb = min(kExQB, nw - g0);
      for (int i = tid; i < kExQB * cap; i += kExThreads) buf[i] = 0ull;
      if (tid < kExQB) { cnt_s[tid] = 0; thr_s[tid] = 0ull; }
      const float* qrow[kExQB];
#pragma unroll
      for (int j = 0; j < kExQB; ++j) qrow[j] = p.Q + int64_t(p.ex_q[w0 + g0 + (j < nqb ? j : 0)]) * p.ldq;
      __syncthreads();
      for (int64_t it = 0; it < iters; ++it) {
        const int64_t r = (it * grid + blockIdx.x) * 8 + warp;
        if (r < p.n) {
          const T* xrow = X + r * int64_t(p.ld);
          double acc[kExQB];
#pragma unroll
          for (int j = 0; j < kExQB; ++j) acc[j] = 0.0;
#pragma unroll 2
          for (int c = lane * 4; c < p.d; c += 128) {
            if (c + 4 <= p.d) {
              const float4 xv = ld4(xrow + c);
#pragma unroll
              for (int j = 0; j < kExQB; ++j) {
                const float4 qv = ld4(qrow[j] + c);
                if (p.metric == 0) {
                  acc[j] = fold1<Red::kDot>(acc[j], qv.x, xv.x);
                  acc[j] = fold1<Red::kDot>(acc[j], qv.y, xv.y);
                  acc[j] = fold1<Red::kDot>(acc[j], qv.z, xv.z);
                  acc[j] = fold1<Red::kDot>(acc[j], qv.w, xv.w);
                } else {
                  acc[j] = fold1<Red::kSqDiff>(acc[j], qv.x, xv.x);
                  acc[j] = fold1<Red::kSqDiff>(acc[j], qv.y, xv.y);
                  acc[j] = fold1<Red::kSqDiff>(acc[j], qv.z, xv.z);
                  acc[j] = fold1<Red::kSqDiff>(acc[j], qv.w, xv.w);
                }
              }
            } else {
              for (int e = c; e < p.d; ++e) {
                const float xe = ld_elem(xrow, e);
#pragma unroll
                for (int j = 0; j < kExQB; ++j)
                  acc[j] = p.metric == 0 ? fold1<Red::kDot>(acc[j], qrow[j][e], xe) : fold1<Red::kSqDiff>(acc[j], qrow[j][e], xe);
              }
            }
          }
          double mine = 0.0;
#pragma unroll
          for (int j = 0; j < kExQB; ++j) {
            const double v = butterfly64(acc[j]);
            if (lane == j) mine = v;
          }
          if (lane < nqb) {
            const uint64_t key = p.metric == 0 ? make_key_max(float(mine), uint32_t(r)) : make_key_min(float(mine), uint32_t(r));
            if (key > thr_s[lane]) buf[lane * cap + atomicAdd(&cnt_s[lane], 1)] = key;   // < cap: at most 8 appends per iteration
          }
        }
        __syncthreads();
        for (int j = 0; j < nqb; ++j) {
          if (cnt_s[j] > cap - 8) {   // uniform: read after the barrier
            block_bitonic_desc(buf + j * cap, cap);
            if (tid == 0) {
              const int c = min(cnt_s[j], p.k);
              cnt_s[j] = c;
              thr_s[j] = c == p.k ? buf[j * cap + p.k - 1] : 0ull;
            }
            for (int i = p.k + tid; i < cap; i += kExThreads) buf[j * cap + i] = 0ull;
            __syncthreads();
          }
        }
      }
      for (int j = 0; j < nqb; ++j) {
        block_bitonic_desc(buf + j * cap, cap);
        uint64_t* out = lists + (size_t(g0 + j) * grid + blockIdx.x) * p.k;
        for (int i = tid; i < p.k; i += kExThreads) out[i] = buf[j * cap + i];
      }
      __syncthreads();
    }
    grid_barrier(p.ex_bar, epoch);
    // ---------------- phase B: merge the per-CTA lists ----------------
    for (int e = blockIdx.x; e < nw; e += grid) {
      const int q = int(p.ex_q[w0 + e]);
      auto cnt_of = [&](int) -> int { return p.k; };
      auto key_of = [&](int slot, int i) -> uint64_t { return lists[(size_t(e) * grid + slot) * p.k + i]; };
      const int got = gather_topk(buf, p.sortn, p.k, grid, cnt_of, key_of, 0u, off_s, &fill_s, nullptr, buf + p.sortn);
      write_topk(p, q, buf, got);
      __syncthreads();
    }
    grid_barrier(p.ex_bar, epoch);
  }
}

}  // namespace b2r
