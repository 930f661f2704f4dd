// common.cuh — shared device helpers of the flat-retrieval kernels.
//
// Everything here is arithmetic that the CPU oracle (oracle/flat_oracle.c) restates
// bit for bit: the order-preserving key packing, and the fixed-order fp64 row
// reductions ("DOT64") that define the scores the engine returns.
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdlib.h>

namespace b2r {

constexpr int kWarp = 32;
constexpr unsigned kFull = 0xffffffffu;

// ---------------------------------------------------------------------------------
// Order-preserving keys.  A candidate is one uint64:
//   high 32 bits: score mapped so that unsigned order == "better" order,
//   low  32 bits: ~id, so that among equal scores the LOWER id is the larger key.
// Sorting keys descending therefore yields (score best-first, id ascending), the
// tie rule BASELINE.json's north_star fixes.  Key 0 means "empty slot".
// ---------------------------------------------------------------------------------
__host__ __device__ __forceinline__ uint32_t f32_to_ordered(float f) {
#ifdef __CUDA_ARCH__
  uint32_t u = __float_as_uint(f);
#else
  union { float f; uint32_t u; } c; c.f = f; uint32_t u = c.u;
#endif
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__host__ __device__ __forceinline__ float ordered_to_f32(uint32_t o) {
  uint32_t u = (o & 0x80000000u) ? (o & 0x7fffffffu) : ~o;
#ifdef __CUDA_ARCH__
  return __uint_as_float(u);
#else
  union { float f; uint32_t u; } c; c.u = u; return c.f;
#endif
}
// "larger is better" key from a score that is maximised (inner product, or the
// phase-1 surrogate q.x - |x|^2/2 of the L2 metric).
__host__ __device__ __forceinline__ uint64_t make_key_max(float score, uint32_t id) {
  return (uint64_t(f32_to_ordered(score)) << 32) | uint64_t(~id);
}
// "larger is better" key from a value that is minimised (squared L2 distance).
__host__ __device__ __forceinline__ uint64_t make_key_min(float dist, uint32_t id) {
  return (uint64_t(~f32_to_ordered(dist)) << 32) | uint64_t(~id);
}
__host__ __device__ __forceinline__ uint32_t key_id(uint64_t k) { return ~uint32_t(k); }
__host__ __device__ __forceinline__ float key_score_max(uint64_t k) { return ordered_to_f32(uint32_t(k >> 32)); }
__host__ __device__ __forceinline__ float key_score_min(uint64_t k) { return ordered_to_f32(~uint32_t(k >> 32)); }

// ---------------------------------------------------------------------------------
// Element access for the two storage types.
// ---------------------------------------------------------------------------------
__device__ __forceinline__ float ld_elem(const float* p, int64_t i) { return p[i]; }
__device__ __forceinline__ float ld_elem(const __nv_bfloat16* p, int64_t i) { return __bfloat162float(p[i]); }

// 4 consecutive elements starting at a 4-aligned element index (row bases are 16B+ aligned).
__device__ __forceinline__ float4 ld4(const float* p) { return *reinterpret_cast<const float4*>(p); }
__device__ __forceinline__ float4 ld4(const __nv_bfloat16* p) {
  uint2 u = *reinterpret_cast<const uint2*>(p);
  float4 r;
  r.x = __uint_as_float(u.x << 16);
  r.y = __uint_as_float(u.x & 0xffff0000u);
  r.z = __uint_as_float(u.y << 16);
  r.w = __uint_as_float(u.y & 0xffff0000u);
  return r;
}

// ---------------------------------------------------------------------------------
// DOT64: the fixed-order fp64 reductions that DEFINE a returned score.
//   element j belongs to lane ((j >> 2) & 31); a lane folds its elements in
//   increasing j with one fma each, starting from +0.0; the 32 lane sums are combined
//   by the xor butterfly 16, 8, 4, 2, 1 (every lane ends with the same value).
// fp32*fp32 is exact in fp64, so the only roundings are the documented additions.
// `a`/`b` must be readable for [0,d); rows padded to a multiple of 4 may be read with
// the vector path when `vec` is true (both pointers 16-byte aligned for fp32).
// ---------------------------------------------------------------------------------
__device__ __forceinline__ double butterfly64(double acc) {
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) acc += __shfl_xor_sync(kFull, acc, off);
  return acc;
}

enum class Red { kDot, kSqDiff };

template <Red R>
__device__ __forceinline__ double fold1(double acc, float a, float b) {
  if (R == Red::kDot) return fma(double(a), double(b), acc);
  double df = double(a) - double(b);
  return fma(df, df, acc);
}

// Warp-collective.  TA/TB in {float, __nv_bfloat16}.  Scalar tail-safe version.
template <Red R, typename TA, typename TB>
__device__ __forceinline__ double warp_reduce64(const TA* a, const TB* b, int d, int lane) {
  double acc = 0.0;
  int c = lane * 4;
  // four chunks per trip: the eight loads are issued before the first fold (rows come from HBM at random
  // addresses; one load in flight per trip made the re-scoring latency-bound).  Same fold order as the plain loop.
  for (; c + 3 * 128 + 4 <= d; c += 4 * 128) {
    float4 va[4], vb[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) { va[u] = ld4(a + c + u * 128); vb[u] = ld4(b + c + u * 128); }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      acc = fold1<R>(acc, va[u].x, vb[u].x);
      acc = fold1<R>(acc, va[u].y, vb[u].y);
      acc = fold1<R>(acc, va[u].z, vb[u].z);
      acc = fold1<R>(acc, va[u].w, vb[u].w);
    }
  }
  for (; c < d; c += 128) {
    if (c + 4 <= d) {
      float4 va = ld4(a + c), vb = ld4(b + c);
      acc = fold1<R>(acc, va.x, vb.x);
      acc = fold1<R>(acc, va.y, vb.y);
      acc = fold1<R>(acc, va.z, vb.z);
      acc = fold1<R>(acc, va.w, vb.w);
    } else {
      for (int j = c; j < d; ++j) acc = fold1<R>(acc, ld_elem(a, j), ld_elem(b, j));
    }
  }
  return butterfly64(acc);
}

// R rows against one query at once: the same per-row arithmetic as warp_reduce64 (a lane folds ITS elements of a
// row in increasing j; which rows share a trip changes nothing), but the R rows' loads are in flight together and
// the query chunk is loaded once — for re-scoring candidate rows scattered over HBM one latency at a time.
template <Red Rd, int R, typename TA, typename TB>
__device__ __forceinline__ void warp_reduce64_rows(const TA* a, const TB* const (&b)[R], int d, int lane, double (&out)[R]) {
  double acc[R];
#pragma unroll
  for (int r = 0; r < R; ++r) acc[r] = 0.0;
#pragma unroll 2
  for (int c = lane * 4; c < d; c += 128) {
    if (c + 4 <= d) {
      const float4 va = ld4(a + c);
      float4 vb[R];
#pragma unroll
      for (int r = 0; r < R; ++r) vb[r] = ld4(b[r] + c);
#pragma unroll
      for (int r = 0; r < R; ++r) {
        acc[r] = fold1<Rd>(acc[r], va.x, vb[r].x);
        acc[r] = fold1<Rd>(acc[r], va.y, vb[r].y);
        acc[r] = fold1<Rd>(acc[r], va.z, vb[r].z);
        acc[r] = fold1<Rd>(acc[r], va.w, vb[r].w);
      }
    } else {
      for (int j = c; j < d; ++j) {
        const float aj = ld_elem(a, j);
#pragma unroll
        for (int r = 0; r < R; ++r) acc[r] = fold1<Rd>(acc[r], aj, ld_elem(b[r], j));
      }
    }
  }
#pragma unroll
  for (int r = 0; r < R; ++r) out[r] = butterfly64(acc[r]);
}

// Same reduction when rows are not known to be 16-byte aligned (caller-owned arrays).
template <Red R>
__device__ __forceinline__ double warp_reduce64_unaligned(const float* a, const float* b, int d, int lane) {
  double acc = 0.0;
  for (int c = lane * 4; c < d; c += 128) {
    int e = min(c + 4, d);
    for (int j = c; j < e; ++j) acc = fold1<R>(acc, a[j], b[j]);
  }
  return butterfly64(acc);
}

// ---------------------------------------------------------------------------------
// Block-wide bitonic sort of n (power of two) uint64 keys in shared memory, descending.
// All threads of the block must call; ends with a __syncthreads().
// ---------------------------------------------------------------------------------
__device__ __forceinline__ void block_bitonic_desc(uint64_t* s, int n) {
  const int tid = threadIdx.x, nt = blockDim.x;
  for (int k = 2; k <= n; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = tid; i < n; i += nt) {
        int p = i ^ j;
        if (p > i) {
          uint64_t a = s[i], b = s[p];
          bool desc = ((i & k) == 0);
          if ((a < b) == desc) { s[i] = b; s[p] = a; }
        }
      }
      __syncthreads();
    }
  }
}

// Warp-wide bitonic sort of 64 keys held two per lane (element e = r*32 + lane for
// register r in {0,1}), descending by element index.  Shuffle-only.
__device__ __forceinline__ void warp_bitonic64_desc(uint64_t& k0, uint64_t& k1, int lane) {
#pragma unroll
  for (int k = 2; k <= 64; k <<= 1) {
#pragma unroll
    for (int j = k >> 1; j > 0; j >>= 1) {
      if (j == 32) {
        // partner is the other register of the same lane; k == 64 here: whole array descending
        uint64_t hi = k0 > k1 ? k0 : k1, lo = k0 > k1 ? k1 : k0;
        k0 = hi; k1 = lo;
      } else {
        uint64_t p0 = __shfl_xor_sync(kFull, k0, j);
        uint64_t p1 = __shfl_xor_sync(kFull, k1, j);
        bool lower = (lane & j) == 0;            // this lane holds the lower index of the pair
        bool desc0 = ((lane & k) == 0);          // k <= 32: direction from the lane bits
        bool desc1 = desc0;
        if (k == 64) { desc0 = true; desc1 = true; }
        // element index of register 1 is lane + 32: bit k (k<=32) of the index equals the lane bit,
        // except k == 32 where register 1 has the bit set.
        if (k == 32) { desc0 = true; desc1 = false; }
        bool take_max0 = (lower == desc0), take_max1 = (lower == desc1);
        k0 = take_max0 ? (k0 > p0 ? k0 : p0) : (k0 < p0 ? k0 : p0);
        k1 = take_max1 ? (k1 > p1 ? k1 : p1) : (k1 < p1 ? k1 : p1);
      }
    }
  }
}

// ---------------------------------------------------------------------------------
// Programmatic dependent launch: the kernels of one search are launched with the
// programmatic-serialization attribute; each starts by letting ITS successor be scheduled
// and then waits for its predecessor to complete (and its writes to be visible) before it
// reads or writes global memory.  The successor's launch latency and prologue thus overlap
// the predecessor's tail.  Both instructions are no-ops in a kernel launched the plain way.
// ---------------------------------------------------------------------------------
__device__ __forceinline__ void pdl_enter() {
  asm volatile("griddepcontrol.launch_dependents;\ngriddepcontrol.wait;" ::: "memory");
}

inline bool pdl_enabled() {
  static const bool on = [] { const char* v = getenv("B2R_PDL"); return !(v && v[0] == '0'); }();
  return on;
}

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), int grid, int block, size_t smem, cudaStream_t st, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(block);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

__device__ __forceinline__ int next_pow2(int v) {
  int p = 1;
  while (p < v) p <<= 1;
  return p;
}

}  // namespace b2r
