// prep.cuh — kernel (a): L2-normalise / cast, one warp per row, 128-bit coalesced loads.
//
// Replaces faiss normalize_L2 / the normalize() inside sklearn cosine_similarity that the
// reference calls at reinforcement_learning_optimization_after_rag.py:70,78,104
// (zero-norm rows stay unchanged, as sklearn's _handle_zeros_in_scale does).
//
// Arithmetic (restated bit for bit by oracle/flat_oracle.c: b2r_oracle_normalize):
//   ss   = DOT64 sum of squares of the fp32 row            (common.cuh)
//   nrm  = (float) sqrt(ss)                                (IEEE fp64 sqrt, RN to fp32)
//   y_j  = nrm == 0 ? x_j : x_j / nrm                      (IEEE fp32 division)
//   bf16 storage then rounds y_j to nearest-even bf16.
// HBM bytes per row: d*4 read + d*s written (s = 4 or 2) + 4 (norm).
#pragma once
#include "common.cuh"

namespace b2r {

__device__ __forceinline__ void st_elem(float* p, int64_t i, float v) { p[i] = v; }
__device__ __forceinline__ void st_elem(__nv_bfloat16* p, int64_t i, float v) { p[i] = __float2bfloat16_rn(v); }
__device__ __forceinline__ float round_store(float v, float*) { return v; }
__device__ __forceinline__ float round_store(float v, __nv_bfloat16*) { return __bfloat162float(__float2bfloat16_rn(v)); }

__device__ __forceinline__ void st4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }
__device__ __forceinline__ void st4(__nv_bfloat16* p, float4 v) {
  __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y), b = __floats2bfloat162_rn(v.z, v.w);
  uint2 u;
  u.x = *reinterpret_cast<uint32_t*>(&a);
  u.y = *reinterpret_cast<uint32_t*>(&b);
  *reinterpret_cast<uint2*>(p) = u;
}

// Rows of `in` ([n,d] fp32, pitch d, possibly unaligned) -> rows of `out` (pitch ld >= d,
// 16B-aligned rows, zero padded to ld).  norm2[r] (optional) = squared norm of the STORED row
// (after normalisation and rounding), as fp32 of its DOT64 value; maxnorm2 (optional) is
// raised to the largest of them with an integer atomicMax (non-negative floats order as ints).
template <typename TOut>
__global__ void __launch_bounds__(256) ingest_rows_kernel(const float* __restrict__ in, int64_t n, int d,
                                                          TOut* __restrict__ out, int ld, int normalize,
                                                          float* __restrict__ norm2, float* __restrict__ maxnorm2) {
  const int lane = threadIdx.x & 31;
  const int64_t warp0 = (int64_t(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = (int64_t(gridDim.x) * blockDim.x) >> 5;
  const bool vec_in = ((d & 3) == 0) && ((reinterpret_cast<uintptr_t>(in) & 15) == 0);
  float wmax = 0.f;
  constexpr int kRegChunks = 8;                     // rows of up to 1024 elements stay in registers: read ONCE
  if (ld <= kRegChunks * 128) {
    const int nch = (ld + 127) / 128;
    for (int64_t r = warp0; r < n; r += nwarps) {
      const float* x = in + r * int64_t(d);
      TOut* y = out + r * int64_t(ld);
      float4 v[kRegChunks];
#pragma unroll
      for (int t = 0; t < kRegChunks; ++t) {
        const int c = lane * 4 + t * 128;
        v[t] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (t < nch && c < ld) {
          if (c + 4 <= d && vec_in) {
            v[t] = __ldcs(reinterpret_cast<const float4*>(x + c));      // streamed: the caller's rows are not read again
          } else {
            if (c + 0 < d) v[t].x = x[c + 0];
            if (c + 1 < d) v[t].y = x[c + 1];
            if (c + 2 < d) v[t].z = x[c + 2];
            if (c + 3 < d) v[t].w = x[c + 3];
          }
        }
      }
      float nrm = 1.f;
      if (normalize) {                               // DOT64 order: a lane folds ITS elements in increasing j (zeros beyond d add nothing)
        double ss = 0.0;
#pragma unroll
        for (int t = 0; t < kRegChunks; ++t) {
          const int c = lane * 4 + t * 128;
          if (t < nch) {
            if (c + 0 < d) ss = fma(double(v[t].x), double(v[t].x), ss);
            if (c + 1 < d) ss = fma(double(v[t].y), double(v[t].y), ss);
            if (c + 2 < d) ss = fma(double(v[t].z), double(v[t].z), ss);
            if (c + 3 < d) ss = fma(double(v[t].w), double(v[t].w), ss);
          }
        }
        nrm = float(sqrt(butterfly64(ss)));
        if (nrm == 0.f) nrm = 1.f;
      }
      double acc = 0.0;
#pragma unroll
      for (int t = 0; t < kRegChunks; ++t) {
        const int c = lane * 4 + t * 128;
        if (t < nch && c < ld) {
          float4 w = v[t];
          if (normalize) { w.x = w.x / nrm; w.y = w.y / nrm; w.z = w.z / nrm; w.w = w.w / nrm; }
          w.x = round_store(w.x, y); w.y = round_store(w.y, y); w.z = round_store(w.z, y); w.w = round_store(w.w, y);
          st4(y + c, w);
          if (c + 0 < d) acc = fma(double(w.x), double(w.x), acc);
          if (c + 1 < d) acc = fma(double(w.y), double(w.y), acc);
          if (c + 2 < d) acc = fma(double(w.z), double(w.z), acc);
          if (c + 3 < d) acc = fma(double(w.w), double(w.w), acc);
        }
      }
      acc = butterfly64(acc);
      const float n2 = float(acc);
      if (norm2 && lane == 0) norm2[r] = n2;
      wmax = fmaxf(wmax, n2);
    }
    if (maxnorm2 && lane == 0 && wmax > 0.f) atomicMax(reinterpret_cast<int*>(maxnorm2), __float_as_int(wmax));
    return;
  }
  for (int64_t r = warp0; r < n; r += nwarps) {      // long rows: norm first, then a second pass (it hits L1/L2)
    const float* x = in + r * int64_t(d);
    TOut* y = out + r * int64_t(ld);
    float nrm = 1.f;
    if (normalize) {
      double ss = vec_in ? warp_reduce64<Red::kDot>(x, x, d, lane) : warp_reduce64_unaligned<Red::kDot>(x, x, d, lane);
      nrm = float(sqrt(ss));
      if (nrm == 0.f) nrm = 1.f;
    }
    // write the stored row, accumulate its squared norm in the same DOT64 order
    double acc = 0.0;
#pragma unroll 4
    for (int c = lane * 4; c < ld; c += 128) {
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (c + 4 <= d && vec_in) {
        v = ld4(x + c);
      } else {
        if (c + 0 < d) v.x = x[c + 0];
        if (c + 1 < d) v.y = x[c + 1];
        if (c + 2 < d) v.z = x[c + 2];
        if (c + 3 < d) v.w = x[c + 3];
      }
      if (normalize) { v.x = v.x / nrm; v.y = v.y / nrm; v.z = v.z / nrm; v.w = v.w / nrm; }
      v.x = round_store(v.x, y); v.y = round_store(v.y, y); v.z = round_store(v.z, y); v.w = round_store(v.w, y);
      st4(y + c, v);
      if (c + 0 < d) acc = fma(double(v.x), double(v.x), acc);
      if (c + 1 < d) acc = fma(double(v.y), double(v.y), acc);
      if (c + 2 < d) acc = fma(double(v.z), double(v.z), acc);
      if (c + 3 < d) acc = fma(double(v.w), double(v.w), acc);
    }
    acc = butterfly64(acc);
    float n2 = float(acc);
    if (norm2 && lane == 0) norm2[r] = n2;
    wmax = fmaxf(wmax, n2);
  }
  if (maxnorm2 && lane == 0 && wmax > 0.f) atomicMax(reinterpret_cast<int*>(maxnorm2), __float_as_int(wmax));
}

// In-place normalisation of caller-owned [n,d] fp32 (pitch d; any alignment).
__global__ void __launch_bounds__(256) normalize_inplace_kernel(float* __restrict__ x, int64_t n, int d) {
  const int lane = threadIdx.x & 31;
  const int64_t warp0 = (int64_t(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = (int64_t(gridDim.x) * blockDim.x) >> 5;
  const bool vec = ((d & 3) == 0) && ((reinterpret_cast<uintptr_t>(x) & 15) == 0);
  for (int64_t r = warp0; r < n; r += nwarps) {
    float* row = x + r * int64_t(d);
    double ss = vec ? warp_reduce64<Red::kDot>(row, row, d, lane) : warp_reduce64_unaligned<Red::kDot>(row, row, d, lane);
    float nrm = float(sqrt(ss));
    if (nrm == 0.f) continue;
    if (vec) {
#pragma unroll 4
      for (int c = lane * 4; c < d; c += 128) {
        float4 v = ld4(row + c);
        v.x = v.x / nrm; v.y = v.y / nrm; v.z = v.z / nrm; v.w = v.w / nrm;
        st4(row + c, v);
      }
    } else {
      for (int j = lane; j < d; j += 32) row[j] = row[j] / nrm;
    }
  }
}

// Batched reward cosine (reference :63-79, :99-104).
//   cos(i,j) = (float)( dot64 / (sqrt(ssa) * sqrt(ssb)) ), a zero norm counts as 1
// one warp per (sample i, doc j); maxcos[i] is reduced by a second tiny kernel-free step:
// each warp raises an ordered-uint atomicMax, decoded by finish_maxcos_kernel.
__global__ void __launch_bounds__(256) cosine_rows_kernel(const float* __restrict__ resp, const float* __restrict__ docs,
                                                          int64_t n, int64_t m, int d, float* __restrict__ cosout,
                                                          uint32_t* __restrict__ maxord) {
  const int lane = threadIdx.x & 31;
  const int64_t warp0 = (int64_t(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = (int64_t(gridDim.x) * blockDim.x) >> 5;
  for (int64_t p = warp0; p < n * m; p += nwarps) {
    int64_t i = p / m;
    const float* a = resp + i * int64_t(d);
    const float* b = docs + p * int64_t(d);
    double dot = warp_reduce64_unaligned<Red::kDot>(a, b, d, lane);
    double sa = warp_reduce64_unaligned<Red::kDot>(a, a, d, lane);
    double sb = warp_reduce64_unaligned<Red::kDot>(b, b, d, lane);
    double na = sqrt(sa), nb = sqrt(sb);
    if (na == 0.0) na = 1.0;
    if (nb == 0.0) nb = 1.0;
    float c = float(dot / (na * nb));
    if (lane == 0) {
      if (cosout) cosout[p] = c;
      if (maxord) atomicMax(maxord + i, f32_to_ordered(c));
    }
  }
}
__global__ void finish_maxcos_kernel(const uint32_t* __restrict__ maxord, int64_t n, float* __restrict__ maxcos) {
  int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i < n) maxcos[i] = maxord[i] == 0u ? 0.0f : ordered_to_f32(maxord[i]);
}

// Query preparation: [nq,d] fp32 (pitch d) -> qf fp32 [nq,ldq] (normalised if asked, rounded to
// bf16 values in fast mode, zero padded), optional bf16 copy qh [nq,ldq], qn2[q] = squared norm
// of the prepared query (fp32 of DOT64).  One warp per query.
__global__ void __launch_bounds__(256) prep_queries_kernel(const float* __restrict__ q, int64_t nq, int d, int ldq,
                                                           int normalize, int round_bf16, float* __restrict__ qf,
                                                           __nv_bfloat16* __restrict__ qh, float* __restrict__ qn2,
                                                           float* __restrict__ qlo /* optional: q - trunc_tf32(q) */,
                                                           uint32_t* __restrict__ thr /* [nq] cleared */,
                                                           unsigned long long* __restrict__ stats /* [4] cleared if asked */,
                                                           uint32_t* __restrict__ fb /* [4] list counters, cleared */) {
  pdl_enter();
  // the per-search thresholds, counters and hand-over lists start at zero: cleared here instead of by memset launches
  if (stats && blockIdx.x == 0 && threadIdx.x < 4) stats[threadIdx.x] = 0ull;
  if (fb && blockIdx.x == 0 && threadIdx.x < 4) fb[threadIdx.x] = 0u;
  const int lane = threadIdx.x & 31;
  const int64_t warp0 = (int64_t(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = (int64_t(gridDim.x) * blockDim.x) >> 5;
  for (int64_t r = warp0; r < nq; r += nwarps) {
    const float* x = q + r * int64_t(d);
    float nrm = 1.f;
    if (normalize) {
      double ss = warp_reduce64_unaligned<Red::kDot>(x, x, d, lane);
      nrm = float(sqrt(ss));
      if (nrm == 0.f) nrm = 1.f;
    }
    double acc = 0.0;
    for (int c = lane * 4; c < ldq; c += 128) {
      float v[4];
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        float e = (c + t < d) ? x[c + t] : 0.f;
        if (normalize) e = e / nrm;
        if (round_bf16) e = __bfloat162float(__float2bfloat16_rn(e));
        v[t] = e;
        if (c + t < d) acc = fma(double(e), double(e), acc);
      }
      st4(qf + r * int64_t(ldq) + c, make_float4(v[0], v[1], v[2], v[3]));
      if (qh) st4(qh + r * int64_t(ldq) + c, make_float4(v[0], v[1], v[2], v[3]));
      if (qlo) {
        float l[4];
#pragma unroll
        for (int t = 0; t < 4; ++t) l[t] = v[t] - __uint_as_float(__float_as_uint(v[t]) & 0xffffe000u);
        st4(qlo + r * int64_t(ldq) + c, make_float4(l[0], l[1], l[2], l[3]));
      }
    }
    acc = butterfly64(acc);
    if (lane == 0) {
      qn2[r] = float(acc);
      if (thr) thr[r] = 0u;
    }
  }
}

}  // namespace b2r

namespace b2r {
// Squared norms of already-stored rows (shard import): norm2[r] = fp32 of DOT64(x_r, x_r); maxnorm2 raised likewise.
template <typename T>
__global__ void __launch_bounds__(256) row_norms_kernel(const T* __restrict__ X, int64_t n, int d, int ld,
                                                        float* __restrict__ norm2, float* __restrict__ maxnorm2) {
  const int lane = threadIdx.x & 31;
  const int64_t warp0 = (int64_t(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = (int64_t(gridDim.x) * blockDim.x) >> 5;
  float wmax = 0.f;
  for (int64_t r = warp0; r < n; r += nwarps) {
    const T* x = X + r * int64_t(ld);
    const float n2 = float(warp_reduce64<Red::kDot>(x, x, d, lane));
    if (norm2 && lane == 0) norm2[r] = n2;
    wmax = fmaxf(wmax, n2);
  }
  if (maxnorm2 && lane == 0 && wmax > 0.f) atomicMax(reinterpret_cast<int*>(maxnorm2), __float_as_int(wmax));
}
// bf16 shadow of stored fp32 rows (exact mode's optional prefilter copy): same pitch, RN-even.
__global__ void __launch_bounds__(256) shadow_rows_kernel(const float* __restrict__ X, int64_t n, int ld,
                                                          __nv_bfloat16* __restrict__ Xs) {
  const int64_t total = n * int64_t(ld / 4);
  for (int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += int64_t(gridDim.x) * blockDim.x)
    st4(Xs + i * 4, ld4(X + i * 4));
}
}  // namespace b2r
