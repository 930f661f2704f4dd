// finalize.cuh — kernel (c): block/grid-level top-k merge, fused with the exact re-scoring.
//
// One CTA per query (1024 threads for few queries, 256 for many):
//   1. gathers the query's candidate keys from every slot (one slot per scoring CTA) — only the valid entries,
//      located through a prefix sum of the per-slot counts — drops everything below the final grid-wide
//      threshold and keeps the best K' — radix select (8 passes of 8-bit digits) when many survive, then a
//      bitonic sort of just those (no candidate list is ever written back to HBM);
//   2. re-scores the K' survivors with the DOT64 arithmetic that defines the returned score
//      (one warp per candidate row, fp64 fixed-order reduction of the stored row);
//   3. sorts them by (fp32 score best-first, id ascending) and writes D/I (or packed keys for
//      the cross-GPU all-gather);
//   4. proves completeness: if fl32(a_min + E) is strictly worse than the k-th returned score,
//      no row that the scoring pass filtered out can belong to the top-k.
//   5. a query whose proof fails is handed on, on the device, to the second pass: its index, its prepared row
//      and the fixed threshold T = s_k - E' (E' = error bound of the range pass) go to a compact list; the
//      range pass (score_umma.cuh, range mode) collects EVERY row whose approximate score reaches T — a
//      superset of the true top-k that needs no gap — and finalize_range_kernel re-scores and orders it.
//      If that set does not fit its buffers (thousands of rows tied with the k-th score) the query goes on
//      to the exhaustive DOT64 scan (score_exact.cuh), which cannot fail.  No host round trip anywhere.
// The same gather merges the all-gathered [G,nq,k] packed lists (merge_packed_kernel) and tightens the thresholds
// between the rounds of a very large shard's scan (tighten_thr_kernel in score_umma.cuh).
#pragma once
#include <float.h>
#include "common.cuh"
#include "prep.cuh"

namespace b2r {

constexpr int kFinThreads = 1024;
constexpr int kFinMaxSlots = 1024;

// A device-side list of queries handed on to a later pass: entry i = (query index, fixed threshold of the pass that
// reads the list, threshold for the pass after that, overflow flag) plus the query's prepared row in compact
// buffers that the pass reads through TMA.  n == nullptr: no such list.
struct FbList {
  uint32_t* n;         // entries so far (cleared by prep_queries_kernel)
  uint32_t* q;
  uint32_t* thr;       // ordered-float threshold for the pass that reads this list
  uint32_t* thr_next;  // ... and for the pass after it (a tighter engine), carried along
  uint32_t* ovf;
  float* qc_f;         // [entry][ldq] prepared rows (fp32), their TF32 low parts, bf16 copies (whichever the pass needs)
  float* qc_lo;
  __nv_bfloat16* qc_h;
};

struct FinalizeParams {
  const uint64_t* cand;    // candidate keys
  int64_t slot_stride;     // elements between consecutive slots
  int64_t q_stride;        // elements between consecutive queries inside a slot
  int nslots;
  int per;                 // capacity (entries) of one (slot, query) list
  const uint32_t* cnt;     // optional valid-entry counts; nullptr = all `per` entries, zeros skipped
  int interleave;          // 0: cand[slot*slot_stride + q*q_stride + i], cnt[slot*nq + q]
                           // 128 (tcgen05 engine): queries are grouped in launches of `qchunk` and tiles of 128;
                           //   cand[(((l*slot_total + slot)*mt_max + m)*per + i)*128 + t], cnt[((l*slot_total + slot)*mt_max + m)*128 + t]
  int mt_max, qchunk, slot_total;   // slot_total: slots per query chunk in the layout (>= nslots scanned)
  const uint32_t* thr;     // optional ordered-float thresholds [nq] (0 = none)
  int nq, kp, k, sortn;
  const void* X; int ld; int d; int storage;
  const float* Q; int ldq; const float* qn2;
  int metric; int64_t n; int64_t row_offset;
  float errfac; const float* maxnorm2;
  float* D; int64_t* I; uint64_t* packed;
  unsigned long long* stats;  // [0] queries unproven by the first pass, [1] candidates seen,
                              // [2] queries sent on to the exhaustive scan, [3] queries left unproven
  // hand-over of unproven queries: finalize_kernel appends to `next`; finalize_range_kernel reads
  // `cur` and appends what does not fit to `next`; whoever has no `next` appends to the exhaustive list
  FbList cur, next;
  uint32_t* ex_n;          // exhaustive list: entries, query indices, grid barrier word (nullptr: count in stats[3] only)
  uint32_t* ex_q;
  uint32_t* ex_bar;
  float errfac_range;      // error factor of the engine that reads `next`, and of the one after it
  float errfac_range2;
  int range_keep;          // finalize_range_kernel: most candidates it takes per query (more -> exhaustive list)
};

// Block-wide radix select: returns the `keep`-th largest of the n distinct keys in s[0..n) (n >= keep >= 1).
// 8-bit digits from the top, O(n) work instead of a full sort; the leading digits on which all keys agree are
// skipped (candidate scores share sign, exponent and the first mantissa bits).  (Merging same-bin lanes with
// __match_any_sync was tried and lost: the match costs more than the serialised shared atomics it saves.)
// hist_s: 256 ints, sel_s: 2 ints.
__device__ __forceinline__ uint64_t block_radix_kth(const uint64_t* s, int n, int keep, int* hist_s, int* sel_s) {
  const int tid = threadIdx.x, nt = blockDim.x, lane = tid & 31;
  // common prefix: OR and AND of all keys (two 64-bit words kept as 4 ints in hist_s[0..3])
  if (tid < 4) hist_s[tid] = (tid & 1) ? -1 : 0;   // [0],[2]: OR halves   [1],[3]: AND halves
  __syncthreads();
  {
    uint64_t o = 0ull, a = ~0ull;
    for (int i = tid; i < n; i += nt) { const uint64_t k = s[i]; o |= k; a &= k; }
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) {
      o |= __shfl_xor_sync(kFull, o, off);
      a &= __shfl_xor_sync(kFull, a, off);
    }
    if (lane == 0) {
      atomicOr(&hist_s[0], int(uint32_t(o)));
      atomicOr(&hist_s[2], int(uint32_t(o >> 32)));
      atomicAnd(&hist_s[1], int(uint32_t(a)));
      atomicAnd(&hist_s[3], int(uint32_t(a >> 32)));
    }
  }
  __syncthreads();
  const uint64_t all_or = (uint64_t(uint32_t(hist_s[2])) << 32) | uint32_t(hist_s[0]);
  const uint64_t all_and = (uint64_t(uint32_t(hist_s[3])) << 32) | uint32_t(hist_s[1]);
  __syncthreads();
  const uint64_t diff = all_or ^ all_and;
  if (diff == 0ull) return all_and;                                  // n == 1 (keys are distinct)
  const int top = (63 - __clzll((long long)diff)) & ~7;              // shift of the first digit that differs
  uint64_t prefix = top == 56 ? 0ull : (all_and & (~0ull << (top + 8)));
  int need = keep;
  for (int shift = top; shift >= 0; shift -= 8) {
    for (int i = tid; i < 256; i += nt) hist_s[i] = 0;
    __syncthreads();
    const uint64_t himask = shift == 56 ? 0ull : (~0ull << (shift + 8));
    for (int i = tid; i < n; i += nt) {
      const uint64_t key = s[i];
      if ((key & himask) == prefix) atomicAdd(&hist_s[int(key >> shift) & 255], 1);
    }
    __syncthreads();
    if (tid < 32) {   // lane l owns digits 255-8l .. 248-8l; find the digit where the count from the top reaches `need`
      int c[8], sum = 0;
#pragma unroll
      for (int j = 0; j < 8; ++j) { c[j] = hist_s[255 - 8 * lane - j]; sum += c[j]; }
      int incl = sum;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        int v = __shfl_up_sync(kFull, incl, o);
        if (lane >= o) incl += v;
      }
      int above = incl - sum;   // keys with a larger digit than any of this lane's
      if (above < need && incl >= need) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          if (above < need && above + c[j] >= need) { sel_s[0] = 255 - 8 * lane - j; sel_s[1] = need - above; above = need; }
          else above += c[j];
        }
      }
    }
    __syncthreads();
    prefix |= uint64_t(sel_s[0]) << shift;
    need = sel_s[1];
    __syncthreads();
  }
  return prefix;
}

// Gathers the keys of `nslots` lists (list s holds cnt_of(s) keys key_of(s, i)), drops zero keys and keys whose
// score half is below thr_hi, and leaves the best `keep` in s[0..keep) sorted descending (s has sortn entries,
// sortn >= keep + 2048).  Returns min(#kept keys, keep).  Block-collective; off_s is [kFinMaxSlots + 1] ints.
template <typename CntF, typename KeyF>
__device__ __forceinline__ int gather_topk(uint64_t* s, int sortn, int keep, int nslots, CntF cnt_of, KeyF key_of,
                                           uint32_t thr_hi, int* off_s, int* fill_s, unsigned long long* seen,
                                           uint64_t* tmp /* [next_pow2(keep)] */) {
  __shared__ int hist_sh[258];   // radix-select histogram + its two result words
  const int tid = threadIdx.x, nt = blockDim.x, lane = tid & 31;
  for (int sl = tid; sl < nslots; sl += nt) off_s[sl + 1] = cnt_of(sl);
  if (tid == 0) { off_s[0] = 0; *fill_s = 0; }
  __syncthreads();
  if (tid < 32) {   // inclusive prefix sum of off_s[1..nslots] by one warp
    const int per_lane = (nslots + 31) / 32;
    const int b = lane * per_lane, e = min(nslots, b + per_lane);
    int sum = 0;
    for (int i = b; i < e; ++i) sum += off_s[i + 1];
    int incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      int v = __shfl_up_sync(kFull, incl, o);
      if (lane >= o) incl += v;
    }
    int run = incl - sum;
    for (int i = b; i < e; ++i) { run += off_s[i + 1]; off_s[i + 1] = run; }
  }
  __syncthreads();
  const int total = off_s[nslots];
  int fill = 0;   // uniform copy of *fill_s
  unsigned long long mine = 0;
  // Gather rounds of nt*U entries: every thread issues its U loads back to back (the entries sit at scattered
  // addresses: one outstanding load per thread and a barrier per round made this loop latency-bound).
  const int U = max(1, min(8, 2048 / nt));
  for (int base = 0; base < total; base += nt * U) {
    if (fill + nt * U > sortn) {   // make room: keep the best `keep` so far (radix select + compaction, O(fill))
      const uint64_t kth = block_radix_kth(s, fill, keep, hist_sh, hist_sh + 256);
      if (tid == 0) *fill_s = 0;
      __syncthreads();
      for (int i = tid; i < fill; i += nt) {
        const uint64_t key = s[i];
        if (key >= kth) {
          const int pos = atomicAdd(fill_s, 1);
          if (pos < keep) tmp[pos] = key;
        }
      }
      __syncthreads();
      for (int i = tid; i < keep; i += nt) s[i] = tmp[i];
      fill = keep;
      if (tid == 0) *fill_s = fill;
      __syncthreads();
    }
    uint64_t keys[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      keys[u] = 0ull;
      const int e = base + u * nt + tid;
      if (u < U && e < total) {
        int lo = 0, hi = nslots;            // largest slot with off_s[slot] <= e
        while (hi - lo > 1) {
          int mid = (lo + hi) >> 1;
          if (off_s[mid] <= e) lo = mid; else hi = mid;
        }
        keys[u] = key_of(lo, e - off_s[lo]);
      }
    }
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const uint64_t key = keys[u];
      if (key != 0ull && uint32_t(key >> 32) >= thr_hi) {
        s[atomicAdd(fill_s, 1)] = key;
        ++mine;
      }
    }
    __syncthreads();
    fill = *fill_s;       // uniform: nobody appends between the two barriers
    __syncthreads();
  }
  if (seen) {   // one global atomic per warp, not per thread
    unsigned long long w = mine;
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) w += __shfl_xor_sync(kFull, w, o);
    if (lane == 0 && w) atomicAdd(seen, w);
  }
  if (fill > 4 * keep && fill > 256) {
    // many more survivors than needed: pick the best `keep` with a radix select, then sort only those
    const uint64_t kth = block_radix_kth(s, fill, keep, hist_sh, hist_sh + 256);
    if (tid == 0) *fill_s = 0;
    __syncthreads();
    for (int i = tid; i < fill; i += nt) {
      const uint64_t key = s[i];
      if (key >= kth) {                                   // exactly `keep` keys: they are distinct
        const int pos = atomicAdd(fill_s, 1);
        if (pos < keep) tmp[pos] = key;
      }
    }
    __syncthreads();
    for (int i = tid; i < keep; i += nt) s[i] = tmp[i];
    fill = keep;
    __syncthreads();
  }
  int nsort = 2;
  while (nsort < fill) nsort <<= 1;
  for (int i = fill + tid; i < nsort; i += nt) s[i] = 0ull;
  __syncthreads();
  block_bitonic_desc(s, nsort);
  return min(fill, keep);
}


// The prepared row of a handed-on query travels to slot e of the list's compact buffers (block-collective).
__device__ __forceinline__ void copy_query_row(const FbList& L, int e, const float* qrow, int ldq, int tid = threadIdx.x,
                                               int nt = blockDim.x) {
  for (int c = tid * 4; c < ldq; c += nt * 4) {
    const float4 v = ld4(qrow + c);
    if (L.qc_f) st4(L.qc_f + int64_t(e) * ldq + c, v);
    if (L.qc_h) st4(L.qc_h + int64_t(e) * ldq + c, v);
    if (L.qc_lo) {
      float4 l;
      l.x = v.x - __uint_as_float(__float_as_uint(v.x) & 0xffffe000u);
      l.y = v.y - __uint_as_float(__float_as_uint(v.y) & 0xffffe000u);
      l.z = v.z - __uint_as_float(__float_as_uint(v.z) & 0xffffe000u);
      l.w = v.w - __uint_as_float(__float_as_uint(v.w) & 0xffffe000u);
      st4(L.qc_lo + int64_t(e) * ldq + c, l);
    }
  }
}

// Rows [0, min(k, ncand)) of the sorted exact keys -> D/I (or packed keys) of query q; the tail is padded.
__device__ __forceinline__ void write_topk(const FinalizeParams& p, int q, const uint64_t* s2, int ncand, int tid = threadIdx.x,
                                           int nt = blockDim.x) {
  for (int j = tid; j < p.k; j += nt) {
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
}

// DOT64 re-scoring of candidates s[0..ncand) (approximate keys: only their ids are used) into exact keys s2[j].
template <typename T>
__device__ __forceinline__ void rescore_keys(const FinalizeParams& p, const float* qrow, const uint64_t* s, int ncand,
                                             uint64_t* s2) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  const T* X = reinterpret_cast<const T*>(p.X);
  const int row_lines = (p.d * int(sizeof(T)) + 127) / 128;     // 128-byte lines of a stored row
  // candidate rows sit at random places of the shard (HBM latency each): the rows this warp scores next are pulled
  // into L2 while it folds the current one
  auto prefetch_row = [&](int j) {
    if (j < ncand) {
      const char* r = reinterpret_cast<const char*>(X + int64_t(key_id(s[j])) * p.ld);
      for (int l = lane; l < row_lines; l += 32) asm volatile("prefetch.global.L2 [%0];" ::"l"(r + size_t(l) * 128));
    }
  };
  if (ncand <= nwarp) {              // a row per warp at most (few queries, big CTAs): nothing to overlap
    if (warp < ncand) {
      const uint32_t id = key_id(s[warp]);
      const T* xrow = X + int64_t(id) * p.ld;
      const uint64_t key = p.metric == 0 ? make_key_max(float(warp_reduce64<Red::kDot>(qrow, xrow, p.d, lane)), id)
                                         : make_key_min(float(warp_reduce64<Red::kSqDiff>(qrow, xrow, p.d, lane)), id);
      if (lane == 0) s2[warp] = key;
    }
    return;
  }
  // four of the warp's rows per trip: their loads are in flight together (same arithmetic per row as one at a time)
  prefetch_row(warp + 4 * nwarp);
  for (int j0 = warp; j0 < ncand; j0 += 4 * nwarp) {
    prefetch_row(j0 + 8 * nwarp);
    uint32_t id[4];
    const T* xrow[4];
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const int j = j0 + r * nwarp;
      id[r] = key_id(s[j < ncand ? j : j0]);
      xrow[r] = X + int64_t(id[r]) * p.ld;
    }
    double out[4];
    if (p.metric == 0) warp_reduce64_rows<Red::kDot, 4>(qrow, xrow, p.d, lane, out);
    else warp_reduce64_rows<Red::kSqDiff, 4>(qrow, xrow, p.d, lane, out);
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const int j = j0 + r * nwarp;
      if (lane == 0 && j < ncand) s2[j] = p.metric == 0 ? make_key_max(float(out[r]), id[r]) : make_key_min(float(out[r]), id[r]);
    }
  }
}

// Completeness proof of query q (one thread): `amin` is the approximate score of the K'-th kept candidate (every
// excluded row scored at most that), `kth` the exact key of the k-th returned row.  A query that cannot be proven is
// counted and appended to the next list (or to the exhaustive list); returns its entry in p.next, else -1.
__device__ __forceinline__ int prove_or_hand_over(const FinalizeParams& p, int q, float amin_f, uint64_t kth) {
  const double amin = double(amin_f);
  const double qn = double(p.qn2[q]), xn = double(*p.maxnorm2);
  const double nn = sqrt(qn * xn);
  const double l2x = 2.4e-7 * (xn + qn + nn);            // L2: rounding of the stored norms and of |q|^2
  bool verified;
  double base, e_fix;     // a row of the true top-k scores >= base - (errfac * nn + e_fix) in a pass with error factor errfac
  if (p.metric == 0) {
    const float sk = key_score_max(kth);
    verified = float(amin + double(p.errfac) * nn) < sk;                 // any excluded row scores <= that bound
    base = double(sk);
    e_fix = 1.2e-7 * nn;                                                 // half an ulp of s_k: keys compare rounded scores
  } else {
    const float dk = key_score_min(kth);
    verified = float(qn - 2.0 * (amin + double(p.errfac) * nn + l2x)) > dk;   // any excluded row is at least that far
    base = 0.5 * (qn - double(dk));
    e_fix = 1.5 * l2x;
  }
  if (verified) return -1;
  atomicAdd(p.stats, 1ull);
  if (p.next.n) {
    const int e = int(atomicAdd(p.next.n, 1u));
    p.next.q[e] = uint32_t(q);
    p.next.thr[e] = f32_to_ordered(__double2float_rd(base - (double(p.errfac_range) * nn + e_fix)));
    p.next.thr_next[e] = f32_to_ordered(__double2float_rd(base - (double(p.errfac_range2) * nn + e_fix)));
    p.next.ovf[e] = 0u;
    return e;
  }
  if (p.ex_n) {
    p.ex_q[atomicAdd(p.ex_n, 1u)] = uint32_t(q);
    atomicAdd(p.stats + 2, 1ull);
  } else {
    atomicAdd(p.stats + 3, 1ull);
  }
  return -1;
}

template <typename T>
__global__ void __launch_bounds__(kFinThreads) finalize_kernel(FinalizeParams p) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  uint64_t* s = reinterpret_cast<uint64_t*>(smem_raw);   // [sortn]
  const int kp2 = next_pow2(p.kp);
  uint64_t* s2 = s + p.sortn;                             // [kp2]
  __shared__ int fill_s;
  __shared__ int off_s[kFinMaxSlots + 1];
  __shared__ int handover_s;                              // -1: proven; >= 0: entry in the range list; -2: exhaustive list
  const int q = blockIdx.x, tid = threadIdx.x;
  pdl_enter();

  const uint32_t thrq = p.thr ? p.thr[q] : 0u;
  const int ql = p.interleave ? q % p.qchunk : 0, lq = p.interleave ? q / p.qchunk : 0;
  auto blk_of = [&](int slot) -> size_t { return (size_t(lq) * p.slot_total + slot) * p.mt_max + (ql >> 7); };
  auto cnt_of = [&](int slot) -> int {
    if (p.interleave) return int(min(p.cnt[blk_of(slot) * 128 + (ql & 127)], uint32_t(p.per)));
    return p.cnt ? int(min(p.cnt[int64_t(slot) * p.nq + q], uint32_t(p.per))) : p.per;
  };
  auto key_of = [&](int slot, int i) -> uint64_t {
    if (p.interleave) return p.cand[(blk_of(slot) * p.per + i) * 128 + (ql & 127)];
    return p.cand[slot * p.slot_stride + q * p.q_stride + i];
  };
  const int ncand = gather_topk(s, p.sortn, p.kp, p.nslots, cnt_of, key_of, thrq, off_s, &fill_s, p.stats ? p.stats + 1 : nullptr, s2);

  // exact re-scoring of the survivors
  const float* qrow = p.Q + int64_t(q) * p.ldq;
  for (int j = tid; j < kp2; j += blockDim.x) s2[j] = 0ull;
  __syncthreads();
  rescore_keys<T>(p, qrow, s, ncand, s2);
  __syncthreads();
  block_bitonic_desc(s2, kp2);
  write_topk(p, q, s2, ncand);

  // completeness proof (thread 0)
  if (tid == 0)
    handover_s = (p.stats && ncand == p.kp && p.n > p.kp)        // the list is full, rows may have been filtered out
                     ? prove_or_hand_over(p, q, key_score_max(s[p.kp - 1]), s2[(p.k < ncand ? p.k : ncand) - 1]) : -1;
  if (!p.next.n) return;     // uniform
  __syncthreads();
  const int e = handover_s;
  if (e < 0) return;
  copy_query_row(p.next, e, qrow, p.ldq);
}

// Second pass of exact mode, per unproven query (entry i of the range list): the range pass left, in the
// candidate scratch, every row whose approximate score reaches the entry's threshold.  They are all re-scored
// (DOT64), ordered, and the top-k overwrites the first pass's rows of D/I.  Entries whose candidate set does not
// fit (a buffer overflowed, or more than range_keep rows) move on to the next list (a pass with a tighter error
// bound) or, when there is none, to the exhaustive list.  CTAs stride over the entries; the launch is a no-op when
// the list is empty.
template <typename T>
__global__ void __launch_bounds__(kFinThreads) finalize_range_kernel(FinalizeParams p) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  uint64_t* s = reinterpret_cast<uint64_t*>(smem_raw);   // [sortn]
  uint64_t* s2 = s + p.sortn;                             // [next_pow2(range_keep)]
  __shared__ int fill_s;
  __shared__ int off_s[kFinMaxSlots + 1];
  const int tid = threadIdx.x;
  __shared__ int next_s;
  pdl_enter();
  const int n1 = int(min(*p.cur.n, uint32_t(p.nq)));
  for (int e = blockIdx.x; e < n1; e += gridDim.x) {
    const int q = int(p.cur.q[e]);
    const int ql = e % p.qchunk, lq = e / p.qchunk;
    auto blk_of = [&](int slot) -> size_t { return (size_t(lq) * p.slot_total + slot) * p.mt_max + (ql >> 7); };
    auto cnt_of = [&](int slot) -> int { return int(min(p.cnt[blk_of(slot) * 128 + (ql & 127)], uint32_t(p.per))); };
    auto key_of = [&](int slot, int i) -> uint64_t { return p.cand[(blk_of(slot) * p.per + i) * 128 + (ql & 127)]; };
    const int got = gather_topk(s, p.sortn, p.range_keep, p.nslots, cnt_of, key_of, 0u, off_s, &fill_s, p.stats + 1, s2);
    const bool fits = p.cur.ovf[e] == 0u && off_s[p.nslots] <= p.range_keep;   // uniform (shared / global, read after barriers)
    __syncthreads();
    if (!fits) {
      // too many rows reach this pass's bound: on to the pass with the tighter bound, or to the exhaustive scan
      if (tid == 0) {
        if (p.next.n) {
          const int e2 = int(atomicAdd(p.next.n, 1u));
          p.next.q[e2] = uint32_t(q);
          p.next.thr[e2] = p.cur.thr_next[e];
          p.next.thr_next[e2] = 0u;
          p.next.ovf[e2] = 0u;
          next_s = e2;
        } else {
          p.ex_q[atomicAdd(p.ex_n, 1u)] = uint32_t(q);
          atomicAdd(p.stats + 2, 1ull);
        }
      }
      if (p.next.n) {
        __syncthreads();
        copy_query_row(p.next, next_s, p.Q + int64_t(q) * p.ldq, p.ldq);
        __syncthreads();
      }
      continue;
    }
    int n2p = 2;
    while (n2p < got) n2p <<= 1;
    for (int j = tid; j < n2p; j += blockDim.x) s2[j] = 0ull;
    __syncthreads();
    rescore_keys<T>(p, p.Q + int64_t(q) * p.ldq, s, got, s2);
    __syncthreads();
    block_bitonic_desc(s2, n2p);
    write_topk(p, q, s2, got);
    __syncthreads();
  }
}

// Merge of G all-gathered packed lists [G][nq][k] -> D/I [nq][k].
__global__ void __launch_bounds__(256) merge_packed_kernel(const uint64_t* __restrict__ packed, int G, int nq, int k,
                                                           int64_t g_stride /* elements between ranks' lists */,
                                                           int sortn, int metric, float* __restrict__ D,
                                                           int64_t* __restrict__ I) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  uint64_t* s = reinterpret_cast<uint64_t*>(smem_raw);
  __shared__ int fill_s;
  __shared__ int off_s[kFinMaxSlots + 1];
  const int q = blockIdx.x;
  pdl_enter();
  auto cnt_of = [&](int) -> int { return k; };
  auto key_of = [&](int g, int i) -> uint64_t { return packed[int64_t(g) * g_stride + int64_t(q) * k + i]; };
  const int got = gather_topk(s, sortn, k, G, cnt_of, key_of, 0u, off_s, &fill_s, nullptr, s + sortn);
  for (int j = threadIdx.x; j < k; j += blockDim.x) {
    const bool have = j < got;
    uint64_t key = have ? s[j] : 0ull;
    float sc = metric == 0 ? (have ? key_score_max(key) : -FLT_MAX) : (have ? key_score_min(key) : FLT_MAX);
    D[int64_t(q) * k + j] = sc;
    I[int64_t(q) * k + j] = have ? int64_t(key_id(key)) : int64_t(-1);
  }
}

}  // namespace b2r
