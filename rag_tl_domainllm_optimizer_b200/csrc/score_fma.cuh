// score_fma.cuh — kernel (b), CUDA-core variant: streaming fp32-FMA scoring fused with a
// threshold-filtered running top-K'.  The [B,N] score matrix never exists in memory.
//
// Used for exact mode at small query batches (true fp32 products, HBM-bound up to ~8 queries
// per pass) and as the engine of last resort for shapes the tcgen05 kernel does not take.
//
// Work split: a CTA (8 warps) walks 32-row tiles of the corpus shard (tile = blockIdx, += grid);
// each warp owns 4 rows of the tile and QB=8 queries held in shared memory.  A lane loads one
// float4 of each of its 4 rows per step (4 x 128-bit coalesced loads in flight, x2 unrolled),
// multiplies with the 8 query float4s (broadcast LDS.128) and keeps 32 partial sums; a
// transposing xor-butterfly (31 shuffles) leaves lane l with the full score of
// (row l>>3, query l&7).  Scores that beat the query's running threshold are appended to the
// CTA's per-query candidate buffer in shared memory; when a buffer fills the CTA sorts it
// (bitonic), keeps the best K' and raises the threshold — published to all CTAs through a
// global atomicMax, so the whole grid filters with the best bound anyone has found.
//
// Algorithmic HBM bytes per pass: rows * d * sizeof(elem) (+ 4*rows for the L2 half-norms).
#pragma once
#include "common.cuh"

namespace b2r {

constexpr int kFmaQB = 8;        // queries per pass
constexpr int kFmaRows = 4;      // rows per warp per tile
constexpr int kFmaWarps = 8;
constexpr int kFmaTile = kFmaRows * kFmaWarps;  // 32 rows per CTA step

struct FmaParams {
  const void* X;         // corpus shard, row pitch ld elements
  int64_t n;             // rows
  int ld;                // row pitch (elements), multiple of 32
  int d;                 // logical dimension (<= ld); rows are zero padded to ld
  const float* Q;        // prepared queries fp32 [nq][ldq]
  int ldq;
  int q0;                // first query of this pass
  int nqb;               // queries in this pass (<= kFmaQB)
  const float* norm2;      // L2 metric: |x|^2 per stored row (fp32), else nullptr
  uint32_t* thr;         // [nq] ordered-float running thresholds shared by all CTAs (0 = none yet)
  uint64_t* cand;        // [grid][nq][kp] candidate keys out (0 = empty)
  int nq;
  int kp;                // K' candidates kept per query
  int cap;               // shared-memory buffer entries per query, power of two >= kp + kFmaTile
};

template <typename T>
__global__ void __launch_bounds__(kFmaWarps * 32, 2) score_fma_kernel(FmaParams p) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float* qs = reinterpret_cast<float*>(smem_raw);                                 // [QB][ld]
  uint64_t* buf = reinterpret_cast<uint64_t*>(qs + size_t(kFmaQB) * p.ld);         // [QB][cap]
  __shared__ int cnt_s[kFmaQB];
  __shared__ float thr_s[kFmaQB];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const T* X = reinterpret_cast<const T*>(p.X);

  // stage the queries of this pass (zero rows beyond nqb)
  for (int i = tid; i < kFmaQB * (p.ld / 4); i += blockDim.x) {
    int b = i / (p.ld / 4), c = (i % (p.ld / 4)) * 4;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (b < p.nqb && c < p.ldq) v = ld4(p.Q + int64_t(p.q0 + b) * p.ldq + c);
    *reinterpret_cast<float4*>(qs + size_t(b) * p.ld + c) = v;
  }
  for (int i = tid; i < kFmaQB * p.cap; i += blockDim.x) buf[i] = 0ull;
  if (tid < kFmaQB) { cnt_s[tid] = 0; thr_s[tid] = -INFINITY; }
  __syncthreads();

  const int64_t ntiles = (p.n + kFmaTile - 1) / kFmaTile;
  const int nchunk = p.ld / 4;  // float4 chunks per row

  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    // refresh the CTA thresholds from the grid-wide ones
    if (tid < p.nqb) {
      uint32_t g = p.thr[p.q0 + tid];
      if (g) thr_s[tid] = fmaxf(thr_s[tid], ordered_to_f32(g));
    }
    const int64_t r0 = tile * kFmaTile + warp * kFmaRows;
    const T* xr[kFmaRows];
#pragma unroll
    for (int r = 0; r < kFmaRows; ++r) {
      int64_t rr = r0 + r < p.n ? r0 + r : p.n - 1;  // clamp: loads stay in bounds, result masked
      xr[r] = X + rr * int64_t(p.ld);
    }
    float acc[kFmaRows * kFmaQB];
#pragma unroll
    for (int i = 0; i < kFmaRows * kFmaQB; ++i) acc[i] = 0.f;

    for (int c = lane; c < nchunk; c += 32) {
      float4 xv[kFmaRows];
#pragma unroll
      for (int r = 0; r < kFmaRows; ++r) xv[r] = ld4(xr[r] + c * 4);
#pragma unroll
      for (int b = 0; b < kFmaQB; ++b) {
        float4 qv = *reinterpret_cast<const float4*>(qs + size_t(b) * p.ld + c * 4);
#pragma unroll
        for (int r = 0; r < kFmaRows; ++r) {
          float a = acc[r * kFmaQB + b];
          a = fmaf(xv[r].x, qv.x, a);
          a = fmaf(xv[r].y, qv.y, a);
          a = fmaf(xv[r].z, qv.z, a);
          a = fmaf(xv[r].w, qv.w, a);
          acc[r * kFmaQB + b] = a;
        }
      }
    }
    // transposing butterfly: 32 values per lane -> lane l holds the sum of value l
#pragma unroll
    for (int off = 16, nn = 16; off >= 1; off >>= 1, nn >>= 1) {
      const bool up = (lane & off) != 0;
#pragma unroll
      for (int i = 0; i < nn; ++i) {
        float send = up ? acc[i] : acc[i + nn];
        float keep = up ? acc[i + nn] : acc[i];
        acc[i] = keep + __shfl_xor_sync(kFull, send, off);
      }
    }
    __syncthreads();  // thr_s refresh visible
    {
      const int r = lane >> 3, b = lane & 7;
      const int64_t row = r0 + r;
      float s = acc[0];
      if (row < p.n && b < p.nqb) {
        if (p.norm2) s -= 0.5f * p.norm2[row];
        if (s >= thr_s[b]) {
          int slot = atomicAdd(&cnt_s[b], 1);
          buf[b * p.cap + slot] = make_key_max(s, uint32_t(row));  // slot < cap by construction
        }
      }
    }
    __syncthreads();
    // compaction of any buffer that could overflow during the next tile
    for (int b = 0; b < p.nqb; ++b) {
      if (cnt_s[b] > p.cap - kFmaTile) {   // uniform: read after the barrier
        block_bitonic_desc(buf + b * p.cap, p.cap);
        if (tid == 0) {
          int c = cnt_s[b] < p.kp ? cnt_s[b] : p.kp;
          cnt_s[b] = c;
          if (c == p.kp) {
            uint64_t kth = buf[b * p.cap + p.kp - 1];
            thr_s[b] = fmaxf(thr_s[b], key_score_max(kth));
            atomicMax(p.thr + p.q0 + b, uint32_t(kth >> 32));
          }
        }
        for (int i = p.kp + tid; i < p.cap; i += blockDim.x) buf[b * p.cap + i] = 0ull;
        __syncthreads();
      }
    }
  }

  // final: best K' of every query of this pass -> global candidate slots of this CTA
  for (int b = 0; b < p.nqb; ++b) {
    block_bitonic_desc(buf + b * p.cap, p.cap);
    uint64_t* out = p.cand + (int64_t(blockIdx.x) * p.nq + (p.q0 + b)) * p.kp;
    for (int i = tid; i < p.kp; i += blockDim.x) out[i] = buf[b * p.cap + i];
    if (tid == 0 && cnt_s[b] >= p.kp) atomicMax(p.thr + p.q0 + b, uint32_t(buf[b * p.cap + p.kp - 1] >> 32));
  }
}

}  // namespace b2r
