// finalize.cuh — kernel (c): block/grid-level top-k merge, fused with the exact re-scoring.
//
// One CTA per query:
//   1. streams the query's candidate keys from every slot (one slot per scoring CTA), drops
//      empties and everything below the final grid-wide threshold, and keeps the best K' with
//      shared-memory bitonic rounds (no candidate list is ever written back to HBM);
//   2. re-scores the K' survivors with the DOT64 arithmetic that defines the returned score
//      (one warp per candidate row, fp64 fixed-order reduction of the stored row);
//   3. sorts them by (fp32 score best-first, id ascending) and writes D/I (or packed keys for
//      the cross-GPU all-gather);
//   4. proves completeness: if fl32(a_min + E) is strictly worse than the k-th returned score,
//      no row that the scoring pass filtered out can belong to the top-k.
// The same kernel body merges the all-gathered [G,nq,k] packed lists (merge_packed_kernel).
#pragma once
#include <float.h>
#include "common.cuh"

namespace b2r {

struct FinalizeParams {
  const uint64_t* cand;    // candidate keys
  int64_t slot_stride;     // elements between consecutive slots
  int64_t q_stride;        // elements between consecutive queries inside a slot
  int nslots;
  int per;                 // entries to scan per (slot, query)
  const uint32_t* cnt;     // optional valid-entry counts; nullptr = scan `per`, skip zeros
  int interleave;          // 0: cand[slot*slot_stride + q*q_stride + i], cnt[slot*nq + q]
                           // 128 (tcgen05 engine): queries are grouped in launches of `qchunk` and tiles of 128;
                           //   cand[(((l*nslots + slot)*mt_max + m)*per + i)*128 + t], cnt[((l*nslots + slot)*mt_max + m)*128 + t]
  int mt_max, qchunk;
  const uint32_t* thr;     // optional ordered-float thresholds [nq] (0 = none)
  int nq, kp, k, sortn;
  const void* X; int ld; int d; int storage;
  const float* Q; int ldq; const float* qn2;
  int metric; int64_t n; int64_t row_offset;
  float errfac; const float* maxnorm2;
  float* D; int64_t* I; uint64_t* packed;
  unsigned long long* stats;  // [0] unverified queries, [1] candidates seen
};

// Keeps the best `keep` of a stream of keys in s[0..keep) (sorted descending), using s[0..sortn).
// fetch(e) returns the e-th raw key or 0.  Returns min(#non-zero keys, keep).  Block-collective.
template <typename Fetch>
__device__ __forceinline__ int stream_topk(uint64_t* s, int sortn, int keep, int64_t total, Fetch fetch,
                                           int* fill_s, unsigned long long* seen) {
  const int tid = threadIdx.x, nt = blockDim.x;
  for (int i = tid; i < sortn; i += nt) s[i] = 0ull;
  if (tid == 0) *fill_s = 0;
  __syncthreads();
  unsigned long long mine = 0;
  int ub = 0;  // uniform upper bound of *fill_s, tracked in registers so the branch never diverges
  for (int64_t base = 0; base < total; base += nt) {
    if (ub + nt > sortn) {
      block_bitonic_desc(s, sortn);
      if (tid == 0) *fill_s = min(*fill_s, keep);
      for (int i = keep + tid; i < sortn; i += nt) s[i] = 0ull;
      __syncthreads();
      ub = keep;
    }
    ub += nt;
    int64_t e = base + tid;
    uint64_t key = e < total ? fetch(e) : 0ull;
    if (key) {
      int pos = atomicAdd(fill_s, 1);
      s[pos] = key;
      ++mine;
    }
    __syncthreads();
  }
  block_bitonic_desc(s, sortn);
  if (seen && mine) atomicAdd(seen, mine);
  int got = min(*fill_s, keep);
  __syncthreads();
  return got;
}

template <typename T>
__global__ void __launch_bounds__(256) finalize_kernel(FinalizeParams p) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  uint64_t* s = reinterpret_cast<uint64_t*>(smem_raw);   // [sortn]
  const int kp2 = next_pow2(p.kp);
  uint64_t* s2 = s + p.sortn;                             // [kp2]
  __shared__ int fill_s;
  const int q = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarp = blockDim.x >> 5;
  const T* X = reinterpret_cast<const T*>(p.X);

  const uint32_t thrq = p.thr ? p.thr[q] : 0u;
  const int ql = p.interleave ? q % p.qchunk : 0, lq = p.interleave ? q / p.qchunk : 0;
  auto fetch = [&](int64_t e) -> uint64_t {
    int slot = int(e / p.per), i = int(e % p.per);
    uint64_t key;
    if (p.interleave) {
      const size_t blk = (size_t(lq) * p.nslots + slot) * p.mt_max + (ql >> 7);
      if (uint32_t(i) >= p.cnt[blk * 128 + (ql & 127)]) return 0ull;
      key = p.cand[(blk * p.per + i) * 128 + (ql & 127)];
    } else {
      if (p.cnt && uint32_t(i) >= p.cnt[int64_t(slot) * p.nq + q]) return 0ull;
      key = p.cand[slot * p.slot_stride + q * p.q_stride + i];
    }
    return (uint32_t(key >> 32) >= thrq) ? key : 0ull;
  };
  const int ncand = stream_topk(s, p.sortn, p.kp, int64_t(p.nslots) * p.per, fetch, &fill_s, p.stats ? p.stats + 1 : nullptr);

  // exact re-scoring of the survivors
  const float* qrow = p.Q + int64_t(q) * p.ldq;
  for (int j = tid; j < kp2; j += blockDim.x) s2[j] = 0ull;
  __syncthreads();
  for (int j = warp; j < ncand; j += nwarp) {
    uint32_t id = key_id(s[j]);
    const T* xrow = X + int64_t(id) * p.ld;
    uint64_t key;
    if (p.metric == 0) {
      float out = float(warp_reduce64<Red::kDot>(qrow, xrow, p.d, lane));
      key = make_key_max(out, id);
    } else {
      float out = float(warp_reduce64<Red::kSqDiff>(qrow, xrow, p.d, lane));
      key = make_key_min(out, id);
    }
    if (lane == 0) s2[j] = key;
  }
  __syncthreads();
  block_bitonic_desc(s2, kp2);

  for (int j = tid; j < p.k; j += blockDim.x) {
    const bool have = j < ncand;
    uint64_t key = have ? s2[j] : 0ull;
    if (p.packed) {
      uint64_t out = 0ull;
      if (have) out = (key & 0xffffffff00000000ull) | uint64_t(~uint32_t(int64_t(key_id(key)) + p.row_offset));
      p.packed[int64_t(q) * p.k + j] = out;
    } else {
      float sc = p.metric == 0 ? (have ? key_score_max(key) : -FLT_MAX) : (have ? key_score_min(key) : FLT_MAX);
      p.D[int64_t(q) * p.k + j] = sc;
      p.I[int64_t(q) * p.k + j] = have ? int64_t(key_id(key)) + p.row_offset : int64_t(-1);
    }
  }

  // completeness proof (thread 0)
  if (tid == 0 && p.stats) {
    bool verified = true;
    if (ncand == p.kp && p.n > p.kp) {           // the list is full, rows may have been filtered out
      const int kk = p.k < ncand ? p.k : ncand;
      double amin = double(key_score_max(s[p.kp - 1]));
      double qn = double(p.qn2[q]), xn = double(*p.maxnorm2);
      double E = double(p.errfac) * sqrt(qn * xn);
      if (p.metric == 0) {
        float bound = float(amin + E);                       // any excluded row scores <= bound
        verified = bound < key_score_max(s2[kk - 1]);
      } else {
        E += 2.4e-7 * (xn + qn + sqrt(qn * xn));
        float dlow = float(qn - 2.0 * (amin + E));           // any excluded row is at distance >= dlow
        verified = dlow > key_score_min(s2[kk - 1]);
      }
    }
    if (!verified) atomicAdd(p.stats, 1ull);
  }
}

// Merge of G all-gathered packed lists [G][nq][k] -> D/I [nq][k].
__global__ void __launch_bounds__(256) merge_packed_kernel(const uint64_t* __restrict__ packed, int G, int nq, int k,
                                                           int sortn, int metric, float* __restrict__ D,
                                                           int64_t* __restrict__ I) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  uint64_t* s = reinterpret_cast<uint64_t*>(smem_raw);
  __shared__ int fill_s;
  const int q = blockIdx.x;
  auto fetch = [&](int64_t e) -> uint64_t {
    int g = int(e / k), i = int(e % k);
    return packed[(int64_t(g) * nq + q) * k + i];
  };
  const int got = stream_topk(s, sortn, k, int64_t(G) * k, fetch, &fill_s, nullptr);
  for (int j = threadIdx.x; j < k; j += blockDim.x) {
    const bool have = j < got;
    uint64_t key = have ? s[j] : 0ull;
    float sc = metric == 0 ? (have ? key_score_max(key) : -FLT_MAX) : (have ? key_score_min(key) : FLT_MAX);
    D[int64_t(q) * k + j] = sc;
    I[int64_t(q) * k + j] = have ? int64_t(key_id(key)) : int64_t(-1);
  }
}

}  // namespace b2r
