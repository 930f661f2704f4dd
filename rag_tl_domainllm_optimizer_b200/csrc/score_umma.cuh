// score_umma.cuh — kernel (b), tensor-core variant: TMA -> shared memory -> tcgen05.mma -> TMEM ->
// threshold-filtered running top-K' in the epilogue.  The [B,N] score matrix never reaches HBM.
//
// One persistent CTA per SM, 6 warps (10 in the 3xTF32 variant):
//   warp 0      TMA producer   : per K-step one 128-row query box (A) and one 256-row corpus box (B),
//                                128 bytes of K per row, SWIZZLE_128B, 4-stage mbarrier ring
//   warp 1      MMA issuer     : one elected lane issues 4 x tcgen05.mma (M=128, N=256, 32 B of K each) per
//                                stage into one of two 256-column TMEM accumulators; tcgen05.commit frees
//                                the smem stage and, after the last K-step, hands the accumulator over
//   warps 2..5  epilogue       : thread = one query (TMEM lane), tcgen05.ld 32 columns at a time, branch-free
//                                pass mask against the query's running threshold (>99.9% of scores fail), then a
//                                warp-uniform walk over the few set bits; survivors are appended to the
//                                (CTA, query) candidate buffer
//   warps 6..9  splitter       : 3xTF32 only: x_lo = x - trunc_tf32(x) on every landed corpus stage
// Work item (t, m) = corpus rows [256t, 256t+256) x queries [128m, 128m+128), t-major, dealt round-robin to the
// CTAs so that the re-reads of a corpus tile by the other query tiles hit L2.  kPair (fp32 rows, >= 2 query tiles):
// two CTAs of a cluster share one M=256 cta_group::2 MMA, each staging half of the corpus tile; whole tiles are
// dealt to the pairs (see the kernel).
//
// Running top-K': threshold + append buffer (cap entries per (CTA, query), interleaved [entry][lane] so that a
// warp's accesses coalesce).  Where the thresholds come from, by launch mode:
//   single    (corpus of at most one tile per CTA) bisection over the 256 scores while they sit in TMEM;
//   sampling  a first launch scores S tiles spread over the shard and keeps only the J best scores of every
//             (tile, query); seed_thr_kernel takes the K'-th largest of them per query; the MAIN launch then
//             covers every row (the sampled ones again) and filters with it;
//   rounds    very large shards: the main scan runs in rounds of 8x growing size, tighten_thr_kernel raises the
//             thresholds to the K'-th best of everything found so far in between;
//   range     fixed thresholds handed over by finalize for the queries its proof could not settle (no compaction:
//             every row that reaches the bound is kept, an overflowing buffer raises a flag).
// When a buffer is nearly full the owning lane selects a new threshold by bisection on the packed keys
// (lane-parallel), keeps [K', kmax] entries and publishes the bound with a global atomicMax, so all CTAs filter
// with the best bound any of them has found.  A published bound is always the K'-th best score of a real subset
// of rows, hence a valid lower bound of the final K'-th best.
//
// Algorithmic HBM bytes per search: rows * ld * sizeof(elem) (the corpus shard, once).
// Algorithmic flops: 2 * nq * rows * d.
#pragma once
#include <cuda.h>
#include <stdlib.h>
#include <algorithm>
#include <string>
#include "common.cuh"
#include "finalize.cuh"

namespace b2r {

constexpr int kUmBM = 128;          // queries per tile (MMA M)
constexpr int kUmBN = 256;          // corpus rows per tile (MMA N)
constexpr int kUmStages = 4;
constexpr int kUmRowBytes = 128;    // bytes of K per row per stage (one swizzle atom)
constexpr int kUmABytes = kUmBM * kUmRowBytes;        // 16 KB
constexpr int kUmBBytes = kUmBN * kUmRowBytes;        // 32 KB
constexpr int kUmStageBytes = kUmABytes + kUmBBytes;  // 48 KB
constexpr int kUmQMax = 2048;       // queries per launch (thresholds / counters live in shared memory)
constexpr int kUmThreads = 192;        // TMA warp, MMA warp, 4 epilogue warps
constexpr int kUmThreadsSplit = 320;   // + 4 splitter warps (3xTF32 variant)
constexpr int kUmSmemState = kUmStages * kUmStageBytes;                 // thr_s follows the stages
constexpr int kUmSmemBytes = kUmSmemState + kUmQMax * 8 + kUmBN * 4 + 256 + 1024;  // + alignment slack

struct UmmaParams {
  int64_t n;          // rows of the shard
  int64_t row0;       // first row covered by this launch (tiles start here; multiple of 256)
  int64_t ntiles;     // corpus tiles covered by this launch
  int nq;             // queries in this launch
  int q0;             // index of the first query of this launch in the whole call
  int mtiles;
  int mt_stride;      // m-tiles per CTA slot in the scratch layout (>= mtiles)
  int num_k;          // K steps per tile
  int kp, cap, kmax;
  int bmax;           // seed bisection stops once K' <= count <= bmax (bmax <= cap - 32)
  const float* norm2; // L2 metric: |x|^2 per stored row
  uint32_t* thr_g;    // [call nq] ordered-float thresholds shared by all CTAs (0 = none yet)
  uint64_t* cand;     // [grid][mtiles][cap][128]
  uint32_t* cnt;      // [grid][mtiles*128]
  uint32_t idesc;
  int slot_base;      // first scratch slot of this launch (seed launch: 0, main launch: #seed CTAs)
  int bisect_first;   // seed launch: find the first tile's threshold by bisection over TMEM
  // range mode (second pass of exact mode): the queries are the entries [q0, q0 + kUmQMax) of the compact list
  // whose length sits in device memory; thresholds (thr_g, indexed like the list) are FIXED; every score that
  // reaches its threshold is kept; a buffer that fills up raises the entry's overflow flag instead of compacting
  int range;
  const uint32_t* nq_dev;   // range mode: length of the whole list
  uint32_t* ovf;            // range mode: [list length] overflow flags
  // sampling launch (large corpora): the launch's `ntiles` tiles are spread evenly over all `all_tiles` tiles of the
  // shard; the epilogue keeps no candidate, only the `seed_j` best scores of every (tile, query), from which
  // seed_thr_kernel derives each query's starting threshold.  The main launch then covers every row.
  int resume;               // later round of a multi-round scan: this CTA's buffers keep what earlier rounds appended
  int seed_topj;
  int seed_j;               // 4 or 8
  int64_t all_tiles;
  float* seedv;             // [call nq][ntiles][seed_j]
#ifdef B2R_UMMA_PROBE
  float* dbg;         // optional [nq][n] dense scores (probe only)
  int dbg_flags;      // probe only: 1 = load the query tile once per CTA, 2 = skip the MMAs, 4 = skip the filter
#endif
};

#ifdef B2R_UMMA_PROBE
#define B2R_DBG_FLAG(p, bit) ((p).dbg_flags & (bit))
#else
#define B2R_DBG_FLAG(p, bit) 0
#endif

// ---------------------------------------------------------------------------------------------------
// PTX wrappers
// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n.reg .pred p;\nelect.sync _|p, 0xffffffff;\nselp.u32 %0, 1, 0, p;\n}" : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  do {
    asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
                 : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
  } while (!ok);
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* tmap, uint32_t bar, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
               ::"r"(dst), "l"(reinterpret_cast<uint64_t>(tmap)), "r"(bar), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tma_load_2d_hint(uint32_t dst, const CUtensorMap* tmap, uint32_t bar, int c0, int c1, uint64_t hint) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%3, %4}], [%2], %5;"
               ::"r"(dst), "l"(reinterpret_cast<uint64_t>(tmap)), "r"(bar), "r"(c0), "r"(c1), "l"(hint) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// ---- CTA pair (cta_group::2): two CTAs of a cluster on the two SMs of a TPC share one M=256 MMA -------------------
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
// shared::cluster address of `local` (a shared::cta address of this CTA) in CTA `cta` of the cluster
__device__ __forceinline__ uint32_t map_to_cta(uint32_t local, uint32_t cta) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local), "r"(cta));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_bar) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_bar) : "memory");
}
__device__ __forceinline__ void mbar_wait_cluster(uint32_t bar, uint32_t parity) {   // arrivals come from the peer CTA too
  uint32_t ok;
  do {
    asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
                 : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
  } while (!ok);
}
// TMA load issued by either CTA of the pair into ITS shared memory, completing on the LEADER's mbarrier
__device__ __forceinline__ void tma_load_2d_pair(uint32_t dst, const CUtensorMap* tmap, uint32_t leader_bar, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
               ::"r"(dst), "l"(reinterpret_cast<uint64_t>(tmap)), "r"(leader_bar), "r"(c0), "r"(c1) : "memory");
}
// commit of the pair's MMAs: one arrival on the barrier at this offset in BOTH CTAs
__device__ __forceinline__ void tc_commit_pair(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(bar), "h"(uint16_t(3)) : "memory");
}
template <bool kTf32>
__device__ __forceinline__ void tc_mma_pair(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
  if (kTf32) {
    asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n}"
                 ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
  } else {
    asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n}"
                 ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
  }
}

template <bool kTf32>
__device__ __forceinline__ void tc_mma(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
  if (kTf32) {
    asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n}"
                 ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
  } else {
    asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}"
                 ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
  }
}
// 32 consecutive fp32 columns of this thread's TMEM lane; returns after the data is in registers.
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&u)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n"
      "tcgen05.wait::ld.sync.aligned;"
      : "=r"(u[0]), "=r"(u[1]), "=r"(u[2]), "=r"(u[3]), "=r"(u[4]), "=r"(u[5]), "=r"(u[6]), "=r"(u[7]), "=r"(u[8]),
        "=r"(u[9]), "=r"(u[10]), "=r"(u[11]), "=r"(u[12]), "=r"(u[13]), "=r"(u[14]), "=r"(u[15]), "=r"(u[16]),
        "=r"(u[17]), "=r"(u[18]), "=r"(u[19]), "=r"(u[20]), "=r"(u[21]), "=r"(u[22]), "=r"(u[23]), "=r"(u[24]),
        "=r"(u[25]), "=r"(u[26]), "=r"(u[27]), "=r"(u[28]), "=r"(u[29]), "=r"(u[30]), "=r"(u[31])
      : "r"(taddr)
      : "memory");
}

// one fp32 column of this thread's TMEM lane (warp-collective, uniform column)
__device__ __forceinline__ uint32_t tmem_ld1(uint32_t taddr) {
  uint32_t v;
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];\ntcgen05.wait::ld.sync.aligned;" : "=r"(v) : "r"(taddr) : "memory");
  return v;
}

// K-major, SWIZZLE_128B shared-memory matrix descriptor (tcgen05 "smem descriptor"):
//   bits [0,14) start address >> 4, [16,30) leading byte offset >> 4 (unused for swizzled K-major: 1),
//   [32,46) stride byte offset >> 4 (8 rows x 128 B = 1024 B), [46,48) version = 1, [61,64) layout 2 = 128B swizzle.
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t saddr) {
  uint64_t d = uint64_t((saddr & 0x3ffffu) >> 4);
  d |= uint64_t(1) << 16;
  d |= uint64_t(1024 >> 4) << 32;
  d |= uint64_t(1) << 46;
  d |= uint64_t(2) << 61;
  return d;
}

// running top-J of a stream of scores, kept sorted (descending) in registers, branch-free
template <int J>
__device__ __forceinline__ void topj_insert(float (&v)[J], float s) {
  v[J - 1] = fmaxf(v[J - 1], s);
#pragma unroll
  for (int i = J - 1; i > 0; --i) {
    const float hi = fmaxf(v[i - 1], v[i]), lo = fminf(v[i - 1], v[i]);
    v[i - 1] = hi;
    v[i] = lo;
  }
}

template <int J, bool kL2>
__device__ __forceinline__ void seed_topj_tile(uint32_t taddr, int ncols, const float* hn_s, float* out) {
  float v[J];
#pragma unroll
  for (int i = 0; i < J; ++i) v[i] = -INFINITY;
  uint32_t u[32];
  for (int c = 0; c < kUmBN / 32; ++c) {
    if (c * 32 >= ncols) break;
    tmem_ld32(taddr + c * 32, u);
#pragma unroll
    for (int i = 0; i < 32; ++i) {
      float s = __uint_as_float(u[i]);
      if (kL2) s -= hn_s[c * 32 + i];
      topj_insert<J>(v, c * 32 + i < ncols ? s : -INFINITY);
    }
  }
  if (out) {
#pragma unroll
    for (int i = 0; i < J; i += 4) *reinterpret_cast<float4*>(out + i) = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
  }
}

// ---------------------------------------------------------------------------------------------------
// lane-parallel compaction of (CTA, query) candidate buffers (entries at buf[e * 128])
// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ void compact_lanes(uint64_t* buf, uint32_t& cnt, float& thr, int kp, int kmax,
                                              uint32_t* thr_g_q, bool need) {
  const uint32_t n = need ? cnt : 0u;
  uint64_t mn = ~0ull, mx = 0ull;
  for (uint32_t e = 0; e < n; ++e) {
    uint64_t k = buf[size_t(e) * 128];
    mn = k < mn ? k : mn;
    mx = k > mx ? k : mx;
  }
  uint64_t lo = mn, hi = mx;
  uint32_t clo = n;
  // invariant: count(key >= lo) == clo >= kp, count(key >= hi + 1) < kp
  while (__any_sync(kFull, need && clo > uint32_t(kmax) && hi > lo)) {
    const bool act = need && clo > uint32_t(kmax) && hi > lo;
    const uint64_t mid = lo + ((hi - lo) >> 1) + 1;   // lo < mid <= hi
    uint32_t c = 0;
    const uint32_t nn = act ? n : 0u;
#pragma unroll 8
    for (uint32_t e = 0; e < nn; ++e) c += (buf[size_t(e) * 128] >= mid) ? 1u : 0u;
    if (act) {
      if (c >= uint32_t(kp)) { lo = mid; clo = c; } else { hi = mid - 1; }
    }
  }
  if (need) {
    uint32_t w = 0;
    for (uint32_t e = 0; e < n; ++e) {
      uint64_t k = buf[size_t(e) * 128];
      if (k >= lo) { buf[size_t(w) * 128] = k; ++w; }
    }
    cnt = w;
    thr = fmaxf(thr, key_score_max(lo));
    atomicMax(thr_g_q, uint32_t(lo >> 32));
  }
}

// ---------------------------------------------------------------------------------------------------
// the kernel
// ---------------------------------------------------------------------------------------------------
template <bool kTf32, bool kL2, bool kSplit, bool kPair>
__global__ void __launch_bounds__(kUmThreadsSplit, 1)
score_umma_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_qlo,
                  const __grid_constant__ CUtensorMap tmap_x, const UmmaParams p) {
  // !kSplit: stage = A 16K | B 32K, 4 stages.
  // kSplit (3xTF32, fp32 rows only): stage = A 16K | A_lo 16K | B 32K | B_lo 32K, 2 stages.  The tensor core
  // truncates fp32 operands to TF32, so A and B as loaded ARE the high parts; A_lo = q - trunc(q) is prepared on
  // the host side of the launch, B_lo = x - trunc(x) is computed from the landed tile by 4 splitter warps
  // (element-wise, so the swizzled layout needs no decoding).  Each K-step then issues q_hi.x_hi + q_lo.x_hi +
  // q_hi.x_lo into the same accumulator: fp32-grade scores (dropped term <= 2^-20) at 3x the MMA work.
  // kPair (cta_group::2, launched as clusters of 2): the pair scores 256 queries (128 per CTA: A, and the TMEM
  // accumulators, stay per CTA) against one 256-row corpus tile of which each CTA stages HALF (B is shared by the
  // two tensor cores): stage = A 16K | B/2 16K, 6 stages.  Only the leader (cluster rank 0) issues MMAs (M = 256);
  // both CTAs' TMA loads complete on the leader's `full` barrier; its commits arrive on both CTAs' `empty` /
  // `tfull` barriers; both CTAs' epilogue warps release an accumulator on the leader's `tempty` barrier.
  static_assert(!(kPair && kSplit), "the CTA-pair variant has no 3xTF32 form");
  constexpr int kStages = kSplit ? 2 : (kPair ? 6 : 4);
  constexpr int kBStage = kPair ? kUmBBytes / 2 : kUmBBytes;   // corpus bytes this CTA stages per K-step
  constexpr int kStageStride = kSplit ? 2 * (kUmABytes + kUmBBytes) : kUmABytes + kBStage;
  constexpr int kAloOff = kUmABytes;
  constexpr int kBOff = kSplit ? 2 * kUmABytes : kUmABytes;
  constexpr int kBloOff = kBOff + kUmBBytes;
  extern __shared__ unsigned char smem_dyn[];
  // 1024-byte alignment for the 128B-swizzled stages
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_dyn) + 1023) & ~uintptr_t(1023));
  float* thr_s = reinterpret_cast<float*>(smem + kUmSmemState);
  uint32_t* cnt_s = reinterpret_cast<uint32_t*>(smem + kUmSmemState + kUmQMax * 4);
  float* hn_s = reinterpret_cast<float*>(smem + kUmSmemState + kUmQMax * 8);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kUmSmemState + kUmQMax * 8 + kUmBN * 4);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 16);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  pdl_enter();
  // queries of this launch: fixed by the host, or (range mode) this launch's share of the device-side list
  int nq = p.nq;
  if (p.range) {
    const int listed = int(*p.nq_dev);
    nq = min(max(listed - p.q0, 0), min(p.nq, kUmQMax));
    if (nq == 0) return;              // uniform over the grid: nothing was left unproven (the normal case)
  }
  const int mtiles = (nq + kUmBM - 1) / kUmBM;
  const uint32_t rank = kPair ? cluster_ctarank() : 0u;
  const uint32_t full0 = smem_u32(bars), empty0 = smem_u32(bars + kStages);
  const uint32_t tfull0 = smem_u32(bars + 2 * kStages), tempty0 = smem_u32(bars + 2 * kStages + 2);
  const uint32_t split0 = smem_u32(bars + 2 * kStages + 4);   // kSplit: "B_lo written" per stage
  const uint32_t stage0 = smem_u32(smem);

  for (int i = threadIdx.x; i < kUmQMax; i += blockDim.x) {
    thr_s[i] = -INFINITY;
    cnt_s[i] = (p.resume && i < mtiles * kUmBM) ? p.cnt[size_t(p.slot_base + blockIdx.x) * p.mt_stride * 128 + i] : 0u;
  }
  if (warp == 0 && lane == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(full0 + 8 * s, 1);
      mbar_init(empty0 + 8 * s, 1);
      if (kSplit) mbar_init(split0 + 8 * s, 4);
    }
    for (int a = 0; a < 2; ++a) { mbar_init(tfull0 + 8 * a, 1); mbar_init(tempty0 + 8 * a, kPair ? 8 : 4); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmap_q)) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmap_x)) : "memory");
  }
  if (warp == 1) {
    if (kPair) {
      asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512u) : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    } else {
      asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512u) : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
  }
  tc_fence_before();
  if (kPair) cluster_sync_all(); else __syncthreads();      // the peer's barriers must be initialised before anything lands on them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // work items = (corpus tile t, query tile m), t-major, dealt round-robin to the CTAs: at any moment the grid
  // works on ~gridDim/mtiles neighbouring corpus tiles, so a tile is fetched from HBM once and its re-reads by the
  // other query tiles hit L2 whatever the tile size (a contiguous split would keep gridDim different tiles live
  // at once: 116 MB for fp32 rows of 768, more than L2 holds)
  const int mitems = kPair ? (mtiles + 1) / 2 : mtiles;      // pair: an item covers two query tiles, one per CTA
  const int64_t nitems = p.ntiles * mitems;
  const int64_t i0 = kPair ? blockIdx.x >> 1 : blockIdx.x, istep = kPair ? gridDim.x >> 1 : gridDim.x;
  // The main launch of the pair kernel deals whole corpus TILES to the pairs instead (pair p takes tiles p, p + pairs,
  // ... and runs all their query-tile pairs back to back): the re-reads of a tile then come from the same two SMs
  // within microseconds and always hit L2.  Measured (ncu dram__bytes_read, 1M x 768 fp32, batch 1024): round-robin
  // items on the pair kernel read the corpus 1.65 times from HBM (four pairs share a tile but drift apart),
  // tiles-per-pair reads it once.
  const bool by_tile = kPair && !p.seed_topj;
  // n-th work item of this CTA (pair): corpus tile t and query-tile group mg; false when there is none
  auto work_item = [&](int64_t n, int64_t& t, int& mg) -> bool {
    if (by_tile) {
      t = i0 + (n / mitems) * istep;
      mg = int(n % mitems);
      return t < p.ntiles;
    }
    const int64_t item = i0 + n * istep;
    t = item / mitems;
    mg = int(item - t * mitems);
    return item < nitems;
  };
  const int kelems = kTf32 ? 32 : 64;   // elements of K per stage

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (elect_one()) {
      uint32_t stage = 0, phase = 0;
      for (int64_t n = 0;; ++n) {
        int64_t t;
        int mg;
        if (!work_item(n, t, mg)) break;
        const int m = kPair ? 2 * mg + int(rank) : mg;
        const int row = int(p.row0 + (p.seed_topj ? (t * p.all_tiles) / p.ntiles : t) * kUmBN) + (kPair ? int(rank) * (kUmBN / 2) : 0);
        for (int k = 0; k < p.num_k; ++k) {
          mbar_wait(empty0 + 8 * stage, phase ^ 1);
          const uint32_t sa = stage0 + stage * kStageStride;
          if (kPair) {
            // both CTAs' boxes complete on the leader's barrier, which expects the bytes of both
            const uint32_t lead = map_to_cta(full0 + 8 * stage, 0);
            if (rank == 0) mbar_expect_tx(full0 + 8 * stage, 2 * (kUmABytes + kBStage));
            tma_load_2d_pair(sa, &tmap_q, lead, k * kelems, m * kUmBM);
            tma_load_2d_pair(sa + kBOff, &tmap_x, lead, k * kelems, row);
            if (++stage == kStages) { stage = 0; phase ^= 1; }
            continue;
          }
          const bool skip_a = B2R_DBG_FLAG(p, 1) && (n > 0 || k >= kStages);
          mbar_expect_tx(full0 + 8 * stage, kUmBBytes + (skip_a ? 0 : (kSplit ? 2 : 1) * kUmABytes));
          if (!skip_a) {
            tma_load_2d(sa, &tmap_q, full0 + 8 * stage, k * kelems, m * kUmBM);
            if (kSplit) tma_load_2d(sa + kAloOff, &tmap_qlo, full0 + 8 * stage, k * kelems, m * kUmBM);
          }
          tma_load_2d(sa + kBOff, &tmap_x, full0 + 8 * stage, k * kelems, row);
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (pair: the leader CTA only) =====================
    uint32_t stage = 0, phase = 0, acc = 0, acc_phase = 0;
    for (int64_t n = 0; !(kPair && rank != 0); ++n) {
      int64_t t_unused;
      int mg_unused;
      if (!work_item(n, t_unused, mg_unused)) break;
      if (kPair) mbar_wait_cluster(tempty0 + 8 * acc, acc_phase ^ 1);
      else mbar_wait(tempty0 + 8 * acc, acc_phase ^ 1);
      tc_fence_after();
      const uint32_t tmem_d = tmem_base + acc * kUmBN;
      for (int k = 0; k < p.num_k; ++k) {
        mbar_wait((kSplit ? split0 : full0) + 8 * stage, phase);   // split done implies the TMA data landed
        tc_fence_after();
        if (elect_one()) {
          const uint32_t sa = stage0 + stage * kStageStride;
          const uint64_t da = make_smem_desc(sa), db = make_smem_desc(sa + kBOff);
#pragma unroll
          for (int j = 0; j < 4; ++j)   // 4 x 32 bytes of K inside the 128-byte swizzle atom
            if (!B2R_DBG_FLAG(p, 2) || (k == 0 && j == 0)) {
              if (kPair) tc_mma_pair<kTf32>(tmem_d, da + uint64_t(2 * j), db + uint64_t(2 * j), p.idesc, uint32_t((k | j) != 0));
              else tc_mma<kTf32>(tmem_d, da + uint64_t(2 * j), db + uint64_t(2 * j), p.idesc, uint32_t((k | j) != 0));
            }
          if (kSplit) {
            const uint64_t dalo = make_smem_desc(sa + kAloOff), dblo = make_smem_desc(sa + kBloOff);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              tc_mma<kTf32>(tmem_d, dalo + uint64_t(2 * j), db + uint64_t(2 * j), p.idesc, 1u);   // q_lo . x_hi
              tc_mma<kTf32>(tmem_d, da + uint64_t(2 * j), dblo + uint64_t(2 * j), p.idesc, 1u);   // q_hi . x_lo
            }
          }
          if (kPair) {
            tc_commit_pair(empty0 + 8 * stage);                       // frees the stage in both CTAs
            if (k == p.num_k - 1) tc_commit_pair(tfull0 + 8 * acc);   // both CTAs' halves of the accumulator are complete
          } else {
            tc_commit(empty0 + 8 * stage);                       // frees the smem stage when the MMAs retire
            if (k == p.num_k - 1) tc_commit(tfull0 + 8 * acc);   // accumulator complete
          }
        }
        __syncwarp();
        if (++stage == kStages) { stage = 0; phase ^= 1; }
      }
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1;
    }
  } else if (kSplit && warp >= 6) {
    // ===================== splitter: B_lo = x - trunc_tf32(x), element-wise on the landed tile =====================
    const int st_tid = (warp - 6) * 32 + lane;   // 0..127
    uint32_t stage = 0, phase = 0;
    for (int64_t n = 0;; ++n) {
      int64_t t_unused;
      int mg_unused;
      if (!work_item(n, t_unused, mg_unused)) break;
      for (int k = 0; k < p.num_k; ++k) {
        mbar_wait(full0 + 8 * stage, phase);
        const uint4* src = reinterpret_cast<const uint4*>(smem + stage * kStageStride + kBOff);
        uint4* dst = reinterpret_cast<uint4*>(smem + stage * kStageStride + kBloOff);
#pragma unroll 4
        for (int i = st_tid; i < kUmBBytes / 16; i += 128) {
          const uint4 v = src[i];
          uint4 o;
          o.x = __float_as_uint(__uint_as_float(v.x) - __uint_as_float(v.x & 0xffffe000u));
          o.y = __float_as_uint(__uint_as_float(v.y) - __uint_as_float(v.y & 0xffffe000u));
          o.z = __float_as_uint(__uint_as_float(v.z) - __uint_as_float(v.z & 0xffffe000u));
          o.w = __float_as_uint(__uint_as_float(v.w) - __uint_as_float(v.w & 0xffffe000u));
          dst[i] = o;
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy writes -> visible to the MMA
        __syncwarp();
        if (lane == 0) mbar_arrive(split0 + 8 * stage);
        if (++stage == kStages) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp < 6) {
    // ===================== epilogue: thread = query =====================
    const int quarter = warp & 3;                 // TMEM lane quarter this warp may read
    const int tq = quarter * 32 + lane;           // query within the tile == TMEM lane
    const int et = (warp - 2) * 32 + lane;        // 0..127 among epilogue threads
    uint32_t acc = 0, acc_phase = 0;
    int64_t t_prev = -1;
    uint32_t seen_m = 0u;   // query-tile groups this CTA has already visited (mitems <= 16)
    for (int64_t n = 0;; ++n) {
      int64_t t;
      int mg;
      if (!work_item(n, t, mg)) break;
      const int m_first = kPair ? 2 * mg + int(rank) : mg;
      const int m_last = m_first + 1;
      const bool first_visit = !((seen_m >> mg) & 1u);   // first time this CTA sees these query tiles
      seen_m |= 1u << mg;
      const int64_t row = p.row0 + (p.seed_topj ? (t * p.all_tiles) / p.ntiles : t) * kUmBN;
      const int ncols = int(p.n - row < kUmBN ? p.n - row : kUmBN);
      if (kL2 && t != t_prev) {
        t_prev = t;
        asm volatile("bar.sync 1, 128;" ::: "memory");          // everyone done with the previous tile's norms
        for (int c = et; c < kUmBN; c += 128) hn_s[c] = c < ncols ? 0.5f * p.norm2[row + c] : 0.f;
        asm volatile("bar.sync 1, 128;" ::: "memory");
      }
      for (int m = m_first; m < m_last; ++m) {
        const int qi = m * kUmBM + tq;
        const bool valid_q = qi < nq;
        const bool warp_has_q = (m * kUmBM + quarter * 32) < nq;
        float thr = -INFINITY;
        uint32_t cnt = 0;
        uint32_t* thr_g_q = p.thr_g + p.q0 + (valid_q ? qi : 0);
        uint32_t g = 0u;
        if (valid_q) {
          thr = thr_s[qi];
          cnt = cnt_s[qi];
          g = *reinterpret_cast<volatile uint32_t*>(thr_g_q);   // issued now, consumed after the wait below
        }
        uint64_t* buf = p.cand + ((size_t(p.slot_base + blockIdx.x) * p.mt_stride + m) * p.cap) * 128 + tq;
        mbar_wait(tfull0 + 8 * acc, acc_phase);
        tc_fence_after();
        if (p.range) thr = valid_q ? ordered_to_f32(g) : INFINITY;   // fixed threshold of the list entry
        else if (g) thr = fmaxf(thr, ordered_to_f32(g));
        const uint32_t taddr = tmem_base + (uint32_t(quarter * 32) << 16) + acc * kUmBN;
        if (p.seed_topj) {
          if (warp_has_q) {
            float* out = valid_q ? p.seedv + (size_t(p.q0 + qi) * p.ntiles + t) * p.seed_j : nullptr;
            if (p.seed_j == 4) seed_topj_tile<4, kL2>(taddr, ncols, hn_s, out);
            else seed_topj_tile<8, kL2>(taddr, ncols, hn_s, out);
          }
        } else if (warp_has_q && !B2R_DBG_FLAG(p, 4)) {
          uint32_t u[32];
          // ---- first tile of this query in this CTA: threshold by bisection over TMEM ----
          if (p.bisect_first && first_visit && ncols >= p.kp) {
            float lo = INFINITY, hi = -INFINITY;
            for (int c = 0; c < kUmBN / 32; ++c) {
              tmem_ld32(taddr + c * 32, u);
#pragma unroll
              for (int i = 0; i < 32; ++i) {
                float s = __uint_as_float(u[i]);
                if (kL2) s -= hn_s[c * 32 + i];
                if (c * 32 + i < ncols) { lo = fminf(lo, s); hi = fmaxf(hi, s); }
              }
            }
            uint32_t clo = uint32_t(ncols);
            for (int it = 0; it < 24; ++it) {
              if (!__any_sync(kFull, clo > uint32_t(p.bmax) && hi > lo)) break;
              const float mid = 0.5f * lo + 0.5f * hi;
              uint32_t cge = 0;
              for (int c = 0; c < kUmBN / 32; ++c) {
                tmem_ld32(taddr + c * 32, u);
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                  float s = __uint_as_float(u[i]);
                  if (kL2) s -= hn_s[c * 32 + i];
                  cge += (c * 32 + i < ncols && s >= mid) ? 1u : 0u;
                }
              }
              if (clo > uint32_t(p.bmax) && hi > lo) {
                if (!(mid > lo && mid < hi)) hi = lo;                       // cannot split further (ties / rounding)
                else if (cge >= uint32_t(p.kp)) { lo = mid; clo = cge; }    // invariant: count(s >= lo) == clo >= kp
                else hi = mid;
              }
            }
            if (valid_q && clo <= uint32_t(p.cap - 32)) {
              thr = fmaxf(thr, lo);
              atomicMax(thr_g_q, f32_to_ordered(lo));
            }
          }
          // ---- filter + append ----
          // Branch-free pass mask per lane (one compare per score), then a warp-uniform walk over the union of
          // the set bits: the passing column is re-read from TMEM with a single-column load (registers cannot be
          // indexed dynamically) and appended by the lanes whose mask has the bit.  >99% of chunks have no bit set.
          for (int c = 0; c < kUmBN / 32; ++c) {
            if (c * 32 >= ncols) break;
            const bool need = valid_q && cnt > uint32_t(p.cap - 32);
            if (p.range) {
              if (need) p.ovf[p.q0 + qi] = 1u;     // too many rows reach the threshold: the exhaustive scan takes over
            } else if (__any_sync(kFull, need)) {
              compact_lanes(buf, cnt, thr, p.kp, p.kmax, thr_g_q, need);
            }
            tmem_ld32(taddr + c * 32, u);
            uint32_t mask = 0u;
#pragma unroll
            for (int i = 0; i < 32; ++i) {
              float s = __uint_as_float(u[i]);
              if (kL2) s -= hn_s[c * 32 + i];
              mask |= (s >= thr) ? (1u << i) : 0u;
            }
            const int left = ncols - c * 32;
            if (left < 32) mask &= (1u << left) - 1u;
            if (!valid_q || (p.range && need)) mask = 0u;
#ifdef B2R_UMMA_PROBE
            if (p.dbg && valid_q) {   // probe only: dense scores
#pragma unroll
              for (int i = 0; i < 32; ++i) {
                float s = __uint_as_float(u[i]);
                if (kL2) s -= hn_s[c * 32 + i];
                if (i < left) p.dbg[size_t(qi) * p.n + row + c * 32 + i] = s;
              }
            }
#endif
            uint32_t any = __reduce_or_sync(kFull, mask);
            if (__popc(any) > 6) {
              // dense chunk (seed tile): walk the statically indexed registers instead of re-reading TMEM
#pragma unroll
              for (int i = 0; i < 32; ++i) {
                if ((mask >> i) & 1u) {
                  float s = __uint_as_float(u[i]);
                  if (kL2) s -= hn_s[c * 32 + i];
                  buf[size_t(cnt) * 128] = make_key_max(s, uint32_t(row + c * 32 + i));
                  ++cnt;
                }
              }
              any = 0u;
            }
            while (any) {
              const int i = __ffs(any) - 1;
              any &= any - 1u;
              const uint32_t v = tmem_ld1(taddr + c * 32 + i);
              if ((mask >> i) & 1u) {
                float s = __uint_as_float(v);
                if (kL2) s -= hn_s[c * 32 + i];
                buf[size_t(cnt) * 128] = make_key_max(s, uint32_t(row + c * 32 + i));
                ++cnt;
              }
            }
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) {
          if (kPair) mbar_arrive_cluster(map_to_cta(tempty0 + 8 * acc, 0));   // the leader's MMA warp waits for both CTAs
          else mbar_arrive(tempty0 + 8 * acc);
        }
        if (valid_q) { thr_s[qi] = thr; cnt_s[qi] = cnt; }
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1;
      }
    }
    // counts of this CTA's buffers
    for (int m = 0; m < (p.seed_topj ? 0 : mtiles); ++m) {
      const int qi = m * kUmBM + tq;
      p.cnt[(size_t(p.slot_base + blockIdx.x) * p.mt_stride + m) * 128 + tq] = qi < nq ? cnt_s[qi] : 0u;
    }
  }

  tc_fence_before();
  if (kPair) cluster_sync_all(); else __syncthreads();      // pair: neither CTA may leave while the other can still reach it
  if (warp == 1) {
    tc_fence_after();
    if (kPair) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
    else asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
  }
}

// ---------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

constexpr int kUmMaxLaunch = 4;   // query chunks of kUmQMax per call (calls are walked in batches of 8192 queries)

// A tensor map is a pure function of (base, shape, pitch, box): encoded once, reused until any of them changes
// (the corpus map changes when rows are added or the storage grows, the query maps when a buffer is re-allocated).
struct TmapSlot {
  bool valid = false, f32 = false;
  const void* base = nullptr;
  uint64_t inner = 0, rows = 0, pitch = 0;
  uint32_t box_rows = 0;
  CUtensorMap map;
};

struct UmmaState {
  PFN_encodeTiled encode = nullptr;
  void* cand = nullptr; size_t cand_bytes = 0;
  void* cnt = nullptr; size_t cnt_bytes = 0;
  void* seedv = nullptr; size_t seedv_bytes = 0;
  bool attr_set[16] = {};
  int max_pairs[4] = {0, 0, 0, 0};     // co-resident CTA pairs of the pair kernel's variants (0 = not asked yet)
  // slots: 0 corpus (first pass), 1 corpus (range pass), 2.. query maps: [pass][lo][launch]
  TmapSlot maps[3 + 3 * 2 * kUmMaxLaunch];
  static constexpr int kMapPairX = 2 + 3 * 2 * kUmMaxLaunch;   // corpus map with 128-row boxes (CTA-pair kernel)
  void release() {
    if (cand) cudaFree(cand);
    if (cnt) cudaFree(cnt);
    if (seedv) cudaFree(seedv);
    cand = cnt = seedv = nullptr;
    cand_bytes = cnt_bytes = seedv_bytes = 0;
  }
  void invalidate() { for (TmapSlot& m : maps) m.valid = false; }
};

struct UmmaCall {
  const void* X; int64_t n; int ld; int d; int storage;
  const void* Qh; const float* Qf; int nq; int kp;
  const float* norm2; uint32_t* thr; int sm_count;
#ifdef B2R_UMMA_PROBE
  float* dbg = nullptr;
  int dbg_flags = 0;
#endif
  bool shadow = false;           // X is the bf16 shadow of fp32 rows: operands carry a 2^-9 rounding each
  bool split = false;            // 3xTF32 (fp32 rows): needs Qlo
  const float* Qlo = nullptr;    // q - trunc_tf32(q), same layout as Qf
  cudaEvent_t ev_main0 = nullptr, ev_main1 = nullptr;   // optional: recorded around the main launch(es)
  int64_t main_rows = 0;                                // out: rows the main launch(es) cover
  // range pass only
  int max_pairs = 0;                    // first pass: co-resident CTA pairs (filled by umma_launch)
  const uint32_t* list_len = nullptr;   // device: entries in the compact list
  uint32_t* ovf = nullptr;              // device: [nq] overflow flags
};

inline bool umma_supported(int storage, int d, int kp, int64_t n) {
  (void)d;
  (void)storage;   // bf16 rows: kind::f16; fp32 rows: kind::tf32 (operands truncated to 10 mantissa bits)
  return kp <= 256 && n > 0 && n < (int64_t(1) << 31);
}

// Error model of the tensor-core pass, relative to |q||x|: products of bf16 (or truncated-TF32) operands are
// exact; every tcgen05.mma folds its K-slice (16 bf16 / 8 tf32 products) into the fp32 accumulator, charged
// here with 2 ulp (2^-23) of the running magnitude per instruction: (d/8 + 8) * 2^-23 for bf16,
// (d/4 + 8) * 2^-23 for tf32 (measured: 2.8e-7 at d=768, tools/umma_probe, 40x below this).  TF32 adds the
// truncation of both operands: |trunc(a) trunc(b) - ab| <= (2^-10 + 2^-10 + 2^-20) |ab|.
// 3xTF32 folds three instructions per K-slice and drops q_lo.x_lo (<= 2^-20) and the TF32 truncation of the two
// low parts (2 * 2^-21).
inline float umma_errfac(bool f32, bool split, bool shadow, int d) {
  return float(split ? (3.0 * double(d) / 4 + 8.0) * 1.1921e-7 + 2.0e-6
               : f32 ? (double(d) / 4 + 8.0) * 1.1921e-7 + 1.96e-3
                     : (double(d) / 8 + 8.0) * 1.1921e-7 + (shadow ? 3.92e-3 : 0.0));
}

inline int umma_get_map(UmmaState& S, int slot, const CUtensorMap** out, bool f32, const void* base, uint64_t inner,
                        uint64_t rows, uint64_t pitch_bytes, uint32_t box_rows, std::string& err) {
  TmapSlot& m = S.maps[slot];
  if (!(m.valid && m.f32 == f32 && m.base == base && m.inner == inner && m.rows == rows && m.pitch == pitch_bytes &&
        m.box_rows == box_rows)) {
    if (!S.encode) {
      void* fn = nullptr;
      cudaDriverEntryPointQueryResult qres;
      cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
      if (e != cudaSuccess || !fn) { err = "cuTensorMapEncodeTiled is not available from this driver"; return -2; }
      S.encode = reinterpret_cast<PFN_encodeTiled>(fn);
    }
    cuuint64_t gdim[2] = {inner, rows};
    cuuint64_t gstride[1] = {pitch_bytes};
    cuuint32_t box[2] = {f32 ? 32u : 64u, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = S.encode(&m.map, f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2,
                          const_cast<void*>(base), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                          CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { m.valid = false; err = "cuTensorMapEncodeTiled failed with code " + std::to_string(int(r)); return -2; }
    m.valid = true; m.f32 = f32; m.base = base; m.inner = inner; m.rows = rows; m.pitch = pitch_bytes; m.box_rows = box_rows;
  }
  *out = &m.map;
  return 0;
}

// Starting thresholds from the sampling launch: thr_g[q] = the K'-th largest of the nvals sampled scores of query q
// (the `seed_j` best of each sampled tile).  Every one of them is the score of a real row, computed exactly as
// the main launch will compute it, so at least K' rows reach it: a valid lower bound of the final K'-th best.
// One warp per query; radix select, one bit per step, over the order-preserving uint image of the scores.
constexpr int kSeedThrWarps = 8;
__global__ void __launch_bounds__(kSeedThrWarps * 32) seed_thr_kernel(const float* __restrict__ seedv, int nq, int nvals,
                                                                       int kp, uint32_t* __restrict__ thr_g) {
  extern __shared__ uint32_t sk_all[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q = blockIdx.x * kSeedThrWarps + warp;
  pdl_enter();
  if (q >= nq) return;
  uint32_t* sk = sk_all + size_t(warp) * nvals;
  const float* v = seedv + size_t(q) * nvals;
  for (int i = lane; i < nvals; i += 32) sk[i] = f32_to_ordered(v[i]);
  __syncwarp();
  uint32_t prefix = 0u;
  for (int bit = 30; bit >= 0; bit -= 2) {           // two bits per step: three independent counts, half the dependent steps
    const uint32_t c1 = prefix | (1u << bit), c2 = prefix | (2u << bit), c3 = prefix | (3u << bit);
    int n1 = 0, n2 = 0, n3 = 0;
    for (int i = lane; i < nvals; i += 32) {
      const uint32_t v = sk[i];
      n1 += v >= c1 ? 1 : 0;
      n2 += v >= c2 ? 1 : 0;
      n3 += v >= c3 ? 1 : 0;
    }
    n1 = __reduce_add_sync(kFull, n1);
    n2 = __reduce_add_sync(kFull, n2);
    n3 = __reduce_add_sync(kFull, n3);
    prefix = n3 >= kp ? c3 : n2 >= kp ? c2 : n1 >= kp ? c1 : prefix;   // the largest prefix that at least K' values reach
  }
  if (lane == 0) thr_g[q] = prefix;     // 0 ("none yet") when fewer than K' values were sampled
}

// Between the rounds of a multi-round scan (very large shards): thr_g[q] is raised to the K'-th best of ALL
// candidates appended so far, over every CTA's buffer — far sharper than any single CTA's own K'-th best, so the
// next (8x larger) round appends only ~8 K' entries per query in total and no buffer ever needs compacting.
// One CTA per query; the candidates stay where they are.
__global__ void __launch_bounds__(256) tighten_thr_kernel(FinalizeParams p, uint32_t* thr_g) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  uint64_t* s = reinterpret_cast<uint64_t*>(smem_raw);
  __shared__ int fill_s;
  __shared__ int off_s[kFinMaxSlots + 1];
  const int q = blockIdx.x;
  pdl_enter();
  const int ql = q % p.qchunk, lq = q / p.qchunk;
  auto blk_of = [&](int slot) -> size_t { return (size_t(lq) * p.slot_total + slot) * p.mt_max + (ql >> 7); };
  auto cnt_of = [&](int slot) -> int { return int(min(p.cnt[blk_of(slot) * 128 + (ql & 127)], uint32_t(p.per))); };
  auto key_of = [&](int slot, int i) -> uint64_t { return p.cand[(blk_of(slot) * p.per + i) * 128 + (ql & 127)]; };
  const int got = gather_topk(s, p.sortn, p.kp, p.nslots, cnt_of, key_of, thr_g[q], off_s, &fill_s, nullptr, s + p.sortn);
  if (threadIdx.x == 0 && got == p.kp) atomicMax(thr_g + q, uint32_t(s[p.kp - 1] >> 32));
}

// Shape of one call's launches and scratch (shared by the first pass and the range pass).
struct UmmaPlan {
  bool f32; int esz, kelems, num_k; int64_t ntiles;
  bool single;           // small corpus: ONE launch (one tile per CTA and query tile, threshold by in-TMEM bisection)
  int tiles_seed;        // sampled tiles of the sampling launch (large corpora)
  int seed_j;
  int grid_seed, grid_main, grid_range, slots, cap, kmax, nlaunch, mt_max;
  bool pair;             // first pass on the CTA-pair kernel (cta_group::2): two query tiles per item
  int nrounds;           // main scan: rounds of growing size (1 unless the shard is very large)
  int64_t round_end[8];  // tile index where round r ends
};

inline int umma_env_int(const char* name, int dflt) {
  const char* v = getenv(name);
  return v && *v ? atoi(v) : dflt;
}

inline UmmaPlan umma_plan(const UmmaCall& c) {
  UmmaPlan P{};
  P.f32 = c.storage == 0;
  P.esz = P.f32 ? 4 : 2;
  P.kelems = P.f32 ? 32 : 64;
  P.num_k = (c.d + P.kelems - 1) / P.kelems;
  P.ntiles = (c.n + kUmBN - 1) / kUmBN;
  const int mt_all = (std::min(c.nq, kUmQMax) + kUmBM - 1) / kUmBM;
  // Small corpus (at most one tile per CTA): a single launch.  With round-robin items CTA b visits query tile
  // (b + j*grid) % mt on its j-th item; a grid size that is 1 modulo mt makes consecutive visits hit different
  // query tiles, so every (CTA, query) gets exactly one tile and its threshold comes from the in-TMEM bisection.
  const int single_cap = mt_all > 1 ? c.sm_count - (c.sm_count - 1) % mt_all : c.sm_count;
  P.single = P.ntiles <= single_cap;
  if (P.single) {
    P.tiles_seed = int(P.ntiles);
    P.grid_seed = int(std::min<int64_t>(c.sm_count, P.ntiles * mt_all));
    if (mt_all > 1 && P.ntiles * mt_all > P.grid_seed) P.grid_seed -= (P.grid_seed - 1) % mt_all;
    P.grid_main = 0;
  } else {
    // Large corpus: a sampling launch over S tiles spread evenly over the shard (their rows are scored again by the
    // main launch: S trades that repeat against the sharpness of the starting threshold, K'/(256 S) of the rows pass)
    // S is chosen in whole waves of the grid: about 1/40 of the scan's items, at least one wave (one query tile: all
    // 148 SMs take one sampled tile each — a sharper threshold for free)
    const int64_t items_want = std::max<int64_t>(1, P.ntiles * mt_all / 40);
    const int waves = int(std::max<int64_t>(1, (items_want + c.sm_count / 2) / c.sm_count));
    // (one query tile is the HBM-bound regime, where the sampling launch costs its tiles' bytes: 1/60 of the shard there)
    const int want = umma_env_int("B2R_SEED_TILES", mt_all == 1 ? int(P.ntiles / 60) : int(int64_t(waves) * c.sm_count / mt_all));
    P.tiles_seed = int(std::min<int64_t>(P.ntiles, std::max(32, std::min(want, c.sm_count))));
    P.seed_j = P.tiles_seed * 4 >= 3 * c.kp ? 4 : 8;
    // tensor-bound fp32 shapes (several query tiles) go to the CTA-pair kernel: each CTA stages half the corpus tile, a
    // third less operand traffic per flop (measured: kind::tf32 2.13 -> 2.05 ms at batch 1024 on 1M x 768; kind::f16
    // loses 6 %, its tensor pipe drops from 89 % to 79 % active, so bf16 rows stay on the single-CTA kernel unless
    // B2R_PAIR=2 asks); an odd tile count wastes half a pair, accepted from 7 tiles up
    const int pair_env = umma_env_int("B2R_PAIR", 1);
    P.pair = !c.split && c.max_pairs > 0 && (pair_env == 2 || (pair_env == 1 && P.f32)) && mt_all >= 2 && (mt_all % 2 == 0 || mt_all >= 7) && c.sm_count >= 2;
    if (P.pair) {
      const int mpairs = (mt_all + 1) / 2, pairs_max = std::min(c.max_pairs, c.sm_count / 2);
      P.grid_seed = 2 * int(std::min<int64_t>(pairs_max, int64_t(P.tiles_seed) * mpairs));
      P.grid_main = 2 * int(std::min<int64_t>(pairs_max, P.ntiles * mpairs));
    } else {
      P.grid_seed = int(std::min<int64_t>(c.sm_count, int64_t(P.tiles_seed) * mt_all));
      P.grid_main = int(std::min<int64_t>(c.sm_count, P.ntiles * mt_all));
    }
  }
  // Very large shards: the sampled threshold lets K'/(256 S) of ALL rows through — tens of thousands of entries per
  // query at 100M rows, and every (CTA, query) buffer would be compacted over and over.  The main scan then runs in
  // rounds of 8x growing size with the thresholds tightened from everything found so far in between (each round
  // appends ~8 K' entries per query in total).
  P.nrounds = 1;
  P.round_end[0] = P.ntiles;
  if (!P.single && P.ntiles > int64_t(umma_env_int("B2R_ROUND_MIN", 64)) * P.tiles_seed) {
    int r = 0;
    for (int64_t e = int64_t(8) * P.tiles_seed; e * 4 < P.ntiles && r < 6; e *= 8) P.round_end[r++] = e;
    P.round_end[r++] = P.ntiles;
    P.nrounds = r;
  }
  P.grid_range = int(std::min<int64_t>(c.sm_count, P.ntiles));
  P.slots = std::max(P.single ? P.grid_seed : P.grid_main, P.grid_range);
  // entries per (CTA, query) buffer: a compaction keeps between K' and kmax = cap/2 - 16 of them, so cap >= 2K' + 32
  P.cap = 128;
  while (P.cap < 2 * c.kp + 32) P.cap <<= 1;
  P.kmax = P.cap / 2 - 16;
  P.nlaunch = (c.nq + kUmQMax - 1) / kUmQMax;
  P.mt_max = mt_all;
  return P;
}

inline int umma_scratch(UmmaState& S, const UmmaPlan& P, std::string& err) {
  const size_t cand_need = size_t(P.nlaunch) * P.slots * P.mt_max * P.cap * 128 * 8;
  const size_t cnt_need = size_t(P.nlaunch) * P.slots * P.mt_max * 128 * 4;
  if (cand_need > S.cand_bytes) {
    if (S.cand) cudaFree(S.cand);
    S.cand = nullptr; S.cand_bytes = 0;
    if (cudaMalloc(&S.cand, cand_need) != cudaSuccess) { cudaGetLastError(); err = "candidate scratch allocation failed"; return -3; }
    S.cand_bytes = cand_need;
  }
  if (cnt_need > S.cnt_bytes) {
    if (S.cnt) cudaFree(S.cnt);
    S.cnt = nullptr; S.cnt_bytes = 0;
    if (cudaMalloc(&S.cnt, cnt_need) != cudaSuccess) { cudaGetLastError(); err = "counter scratch allocation failed"; return -3; }
    S.cnt_bytes = cnt_need;
  }
  return 0;
}

template <bool TF, bool L2, bool SP, bool PR>
inline cudaError_t umma_launch_variant(UmmaState& S, int vi, int grid, const CUtensorMap& tmq, const CUtensorMap& tmqlo,
                                       const CUtensorMap& tmx, const UmmaParams& p, cudaStream_t st) {
  cudaError_t e = cudaSuccess;
  if (!S.attr_set[vi]) {
    e = cudaFuncSetAttribute(score_umma_kernel<TF, L2, SP, PR>, cudaFuncAttributeMaxDynamicSharedMemorySize, kUmSmemBytes);
    if (e != cudaSuccess) return e;
    S.attr_set[vi] = true;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(SP ? kUmThreadsSplit : kUmThreads);
  cfg.dynamicSmemBytes = kUmSmemBytes;
  cfg.stream = st;
  cudaLaunchAttribute at[2];
  int na = 0;
  if (PR) {                      // the two CTAs of a pair must sit on the two SMs of one TPC: a cluster of 2
    at[na].id = cudaLaunchAttributeClusterDimension;
    at[na].val.clusterDim.x = 2; at[na].val.clusterDim.y = 1; at[na].val.clusterDim.z = 1;
    ++na;
  }
  if (pdl_enabled()) {
    at[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  cfg.attrs = at;
  cfg.numAttrs = na;
  return cudaLaunchKernelEx(&cfg, score_umma_kernel<TF, L2, SP, PR>, tmq, tmqlo, tmx, p);
}

inline cudaError_t umma_kernel_launch(UmmaState& S, const UmmaCall& c, bool f32, int grid, const CUtensorMap& tmq,
                                      const CUtensorMap& tmqlo, const CUtensorMap& tmx, const UmmaParams& p, cudaStream_t st,
                                      bool pair = false) {
  const int variant = (f32 ? 2 : 0) + (c.norm2 ? 1 : 0);
  cudaError_t e;
  if (pair) {
    const int vi = 8 + variant;
    if (f32) e = c.norm2 ? umma_launch_variant<true, true, false, true>(S, vi, grid, tmq, tmqlo, tmx, p, st)
                         : umma_launch_variant<true, false, false, true>(S, vi, grid, tmq, tmqlo, tmx, p, st);
    else     e = c.norm2 ? umma_launch_variant<false, true, false, true>(S, vi, grid, tmq, tmqlo, tmx, p, st)
                         : umma_launch_variant<false, false, false, true>(S, vi, grid, tmq, tmqlo, tmx, p, st);
  } else if (c.split) {   // 3xTF32: fp32 rows only
    const int vi = variant * 2 + 1;
    e = c.norm2 ? umma_launch_variant<true, true, true, false>(S, vi, grid, tmq, tmqlo, tmx, p, st)
                : umma_launch_variant<true, false, true, false>(S, vi, grid, tmq, tmqlo, tmx, p, st);
  } else {
    const int vi = variant * 2;
    if (f32) e = c.norm2 ? umma_launch_variant<true, true, false, false>(S, vi, grid, tmq, tmqlo, tmx, p, st)
                         : umma_launch_variant<true, false, false, false>(S, vi, grid, tmq, tmqlo, tmx, p, st);
    else     e = c.norm2 ? umma_launch_variant<false, true, false, false>(S, vi, grid, tmq, tmqlo, tmx, p, st)
                         : umma_launch_variant<false, false, false, false>(S, vi, grid, tmq, tmqlo, tmx, p, st);
  }
  if (e == cudaSuccess) e = cudaGetLastError();
  return e;
}

// How many CTA pairs of the pair kernel can be resident at once (asked once per variant).
inline int umma_max_pairs(UmmaState& S, bool f32, bool l2) {
  const int variant = (f32 ? 2 : 0) + (l2 ? 1 : 0);
  if (S.max_pairs[variant] > 0) return S.max_pairs[variant];
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(2);
  cfg.blockDim = dim3(kUmThreads);
  cfg.dynamicSmemBytes = kUmSmemBytes;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  int n = 0;
  cudaError_t e;
#define B2R_PAIR_OCC(TF, L2)                                                                                           \
  do {                                                                                                                  \
    e = cudaFuncSetAttribute(score_umma_kernel<TF, L2, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kUmSmemBytes); \
    if (e == cudaSuccess) e = cudaOccupancyMaxActiveClusters(&n, score_umma_kernel<TF, L2, false, true>, &cfg);         \
  } while (0)
  if (f32) { if (l2) B2R_PAIR_OCC(true, true); else B2R_PAIR_OCC(true, false); }
  else     { if (l2) B2R_PAIR_OCC(false, true); else B2R_PAIR_OCC(false, false); }
#undef B2R_PAIR_OCC
  if (e != cudaSuccess) { cudaGetLastError(); n = 0; }
  S.max_pairs[variant] = n > 0 ? n : -1;
  return S.max_pairs[variant];
}

inline uint32_t umma_idesc(bool f32, bool pair = false) {   // pair: M = 256 over the two CTAs
  return (1u << 4) | ((f32 ? 2u : 1u) << 7) | ((f32 ? 2u : 1u) << 10) | (uint32_t(kUmBN >> 3) << 17) |
         (uint32_t((pair ? 2 * kUmBM : kUmBM) >> 4) << 24);
}

// Query maps of launch l.  Rows are padded to whole 128-query boxes (the buffers are sized for it; rows >= nq only
// feed ignored outputs): boxes that stay inside the tensor avoid TMA's out-of-bounds handling on every load.
inline int umma_query_maps(UmmaState& S, const UmmaCall& c, const UmmaPlan& P, int pass, int l, int nq,
                           const CUtensorMap** tmq, const CUtensorMap** tmqlo, std::string& err) {
  const int q0 = l * kUmQMax;
  const char* qbase = P.f32 ? reinterpret_cast<const char*>(c.Qf) : reinterpret_cast<const char*>(c.Qh);
  const uint64_t nq_box = uint64_t((nq + 2 * kUmBM - 1) / (2 * kUmBM)) * (2 * kUmBM);
  const int base_slot = 2 + pass * 2 * kUmMaxLaunch;
  int rc = umma_get_map(S, base_slot + l, tmq, P.f32, qbase + size_t(q0) * c.ld * P.esz, uint64_t(c.ld), nq_box,
                        uint64_t(c.ld) * P.esz, kUmBM, err);
  if (rc) return rc;
  const char* qlo = c.split ? reinterpret_cast<const char*>(c.Qlo) : qbase;   // unused map when not split
  return umma_get_map(S, base_slot + kUmMaxLaunch + l, tmqlo, P.f32, qlo + size_t(q0) * c.ld * P.esz, uint64_t(c.ld), nq_box,
                      uint64_t(c.ld) * P.esz, kUmBM, err);
}

inline void umma_fill_finalize(const UmmaState& S, const UmmaPlan& P, int nslots, int slot_total, FinalizeParams& fp) {
  // layout of the scratch for the merge kernels: slot = CTA of either launch, interleaved [m][entry][128]
  fp.cand = reinterpret_cast<const uint64_t*>(S.cand);
  fp.cnt = reinterpret_cast<const uint32_t*>(S.cnt);
  fp.nslots = nslots;
  fp.slot_total = slot_total;
  fp.per = P.cap;
  fp.interleave = 128;
  fp.mt_max = P.mt_max;
  fp.qchunk = kUmQMax;
  fp.slot_stride = 0;
  fp.q_stride = 0;
}

// First pass.  Small corpus: one launch (in-TMEM bisection).  Large corpus: sampling launch -> seed_thr_kernel ->
// main launch over every row (the timers bracket the main launch).
inline int umma_launch(UmmaState& S, UmmaCall& c, cudaStream_t st, FinalizeParams& fp, int* launches, std::string& err) {
  if (c.nq > kUmMaxLaunch * kUmQMax) { err = "too many queries for one tcgen05 call"; return -5; }
  c.max_pairs = umma_max_pairs(S, c.storage == 0, c.norm2 != nullptr);
  const UmmaPlan P = umma_plan(c);       // (no room for clusters, max_pairs < 1: the plan stays on the single-CTA kernel)
  int rc = umma_scratch(S, P, err);
  if (rc) return rc;
  const CUtensorMap* tmx = nullptr;
  if (P.pair) rc = umma_get_map(S, UmmaState::kMapPairX, &tmx, P.f32, c.X, uint64_t(c.ld), uint64_t(c.n), uint64_t(c.ld) * P.esz, kUmBN / 2, err);
  else rc = umma_get_map(S, 0, &tmx, P.f32, c.X, uint64_t(c.ld), uint64_t(c.n), uint64_t(c.ld) * P.esz, kUmBN, err);
  if (rc) return rc;
  const int slots = P.single ? P.grid_seed : P.grid_main;
  umma_fill_finalize(S, P, slots, slots, fp);
  fp.errfac = umma_errfac(P.f32, c.split, c.shadow, c.d);

  const CUtensorMap* tmq[kUmMaxLaunch];
  const CUtensorMap* tmqlo[kUmMaxLaunch];
  UmmaParams pp[kUmMaxLaunch];
  for (int l = 0; l < P.nlaunch; ++l) {
    const int q0 = l * kUmQMax;
    const int nq = std::min(kUmQMax, c.nq - q0);
    rc = umma_query_maps(S, c, P, 0, l, nq, &tmq[l], &tmqlo[l], err);
    if (rc) return rc;
    UmmaParams& p = pp[l];
    p = UmmaParams{};
    p.n = c.n; p.nq = nq; p.q0 = q0; p.mtiles = (nq + kUmBM - 1) / kUmBM; p.mt_stride = P.mt_max; p.num_k = P.num_k;
    p.kp = c.kp; p.cap = P.cap; p.kmax = P.kmax; p.bmax = std::min(P.cap - 40, std::max(2 * c.kp, c.kp + 56)); p.norm2 = c.norm2; p.thr_g = c.thr;
    p.cand = reinterpret_cast<uint64_t*>(S.cand) + size_t(l) * slots * P.mt_max * P.cap * 128;
    p.cnt = reinterpret_cast<uint32_t*>(S.cnt) + size_t(l) * slots * P.mt_max * 128;
    p.idesc = umma_idesc(P.f32, P.pair);
    p.all_tiles = P.ntiles;
#ifdef B2R_UMMA_PROBE
    p.dbg = c.dbg; p.dbg_flags = c.dbg_flags;
#endif
  }
  if (P.single) {
    for (int l = 0; l < P.nlaunch; ++l) {
      UmmaParams& p = pp[l];
      p.row0 = 0; p.ntiles = P.ntiles; p.slot_base = 0; p.bisect_first = 1;
      cudaError_t e = umma_kernel_launch(S, c, P.f32, P.grid_seed, *tmq[l], *tmqlo[l], *tmx, p, st);
      if (e != cudaSuccess) { err = std::string("tcgen05 launch failed: ") + cudaGetErrorString(e); return -2; }
      ++*launches;
    }
    c.main_rows = 0;      // nothing to time separately
    return 0;
  }
  // sampling launch + thresholds
  const int nvals = P.tiles_seed * P.seed_j;
  const size_t need = size_t(c.nq) * nvals * sizeof(float);
  if (need > S.seedv_bytes) {
    if (S.seedv) cudaFree(S.seedv);
    S.seedv = nullptr; S.seedv_bytes = 0;
    if (cudaMalloc(&S.seedv, need) != cudaSuccess) { cudaGetLastError(); err = "sample scratch allocation failed"; return -3; }
    S.seedv_bytes = need;
  }
  for (int l = 0; l < P.nlaunch; ++l) {
    UmmaParams p = pp[l];
    p.row0 = 0; p.ntiles = P.tiles_seed; p.slot_base = 0; p.bisect_first = 0;
    p.seed_topj = 1; p.seed_j = P.seed_j; p.seedv = static_cast<float*>(S.seedv);
    cudaError_t e = umma_kernel_launch(S, c, P.f32, P.grid_seed, *tmq[l], *tmqlo[l], *tmx, p, st, P.pair);
    if (e != cudaSuccess) { err = std::string("tcgen05 sampling launch failed: ") + cudaGetErrorString(e); return -2; }
    ++*launches;
  }
  {
    const size_t smem = size_t(kSeedThrWarps) * nvals * sizeof(uint32_t);
    cudaError_t e = launch_pdl(seed_thr_kernel, (c.nq + kSeedThrWarps - 1) / kSeedThrWarps, kSeedThrWarps * 32, smem, st,
                               static_cast<const float*>(S.seedv), c.nq, nvals, c.kp, c.thr);
    if (e == cudaSuccess) e = cudaGetLastError();
    if (e != cudaSuccess) { err = std::string("seed_thr launch failed: ") + cudaGetErrorString(e); return -2; }
    ++*launches;
  }
  if (c.ev_main0) cudaEventRecord(c.ev_main0, st);
  c.main_rows = c.n;
  int64_t tile0 = 0;
  for (int r = 0; r < P.nrounds; ++r) {
    if (r > 0) {      // thresholds from everything the earlier rounds found
      FinalizeParams tp = fp;
      tp.kp = c.kp;
      tp.sortn = 8192;
      if (!S.attr_set[12]) {
        cudaError_t ea = cudaFuncSetAttribute(tighten_thr_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (8192 + 512) * 8);
        if (ea != cudaSuccess) { err = std::string("tighten_thr attribute failed: ") + cudaGetErrorString(ea); return -2; }
        S.attr_set[12] = true;
      }
      cudaError_t e = launch_pdl(tighten_thr_kernel, c.nq, 256, size_t(tp.sortn + 512) * 8, st, tp, c.thr);
      if (e == cudaSuccess) e = cudaGetLastError();
      if (e != cudaSuccess) { err = std::string("tighten_thr launch failed: ") + cudaGetErrorString(e); return -2; }
      ++*launches;
    }
    for (int l = 0; l < P.nlaunch; ++l) {
      UmmaParams& p = pp[l];
      p.row0 = tile0 * kUmBN; p.ntiles = P.round_end[r] - tile0; p.slot_base = 0; p.bisect_first = 0; p.resume = r > 0;
      cudaError_t e = umma_kernel_launch(S, c, P.f32, P.grid_main, *tmq[l], *tmqlo[l], *tmx, p, st, P.pair);
      if (e != cudaSuccess) { err = std::string("tcgen05 main launch failed: ") + cudaGetErrorString(e); return -2; }
      ++*launches;
    }
    tile0 = P.round_end[r];
  }
  if (c.ev_main1) cudaEventRecord(c.ev_main1, st);
  return 0;
}

// Range pass (second pass of exact mode): one launch per query chunk over ALL rows, with the fixed thresholds of
// the compact list; `c` describes the compact query buffers and the engine of this pass (fp32 rows: 3xTF32),
// `first` the first pass's call (its plan sized the scratch this pass reuses).
inline int umma_launch_range(UmmaState& S, UmmaCall& c, const UmmaCall& first, int pass /* 1 or 2 */, cudaStream_t st,
                             FinalizeParams& fp, int* launches, std::string& err) {
  UmmaPlan P = umma_plan(first);       // same scratch geometry (cap, mt_max, nlaunch, slots) as the first pass
  const bool f32 = c.storage == 0;
  P.f32 = f32; P.esz = f32 ? 4 : 2; P.kelems = f32 ? 32 : 64; P.num_k = (c.d + P.kelems - 1) / P.kelems;
  const CUtensorMap* tmx = nullptr;
  int rc = umma_get_map(S, 1, &tmx, f32, c.X, uint64_t(c.ld), uint64_t(c.n), uint64_t(c.ld) * P.esz, kUmBN, err);
  if (rc) return rc;
  umma_fill_finalize(S, P, P.grid_range, P.grid_range, fp);
  for (int l = 0; l < P.nlaunch; ++l) {
    const int q0 = l * kUmQMax;
    const int nq = std::min(kUmQMax, c.nq - q0);
    const CUtensorMap* tmq = nullptr;
    const CUtensorMap* tmqlo = nullptr;
    rc = umma_query_maps(S, c, P, pass, l, nq, &tmq, &tmqlo, err);
    if (rc) return rc;
    UmmaParams p{};
    p.n = c.n; p.nq = nq; p.q0 = q0; p.mtiles = (nq + kUmBM - 1) / kUmBM; p.mt_stride = P.mt_max; p.num_k = P.num_k;
    p.kp = c.kp; p.cap = P.cap; p.kmax = P.kmax; p.bmax = 0; p.norm2 = c.norm2; p.thr_g = c.thr;
    p.cand = reinterpret_cast<uint64_t*>(S.cand) + size_t(l) * P.grid_range * P.mt_max * P.cap * 128;
    p.cnt = reinterpret_cast<uint32_t*>(S.cnt) + size_t(l) * P.grid_range * P.mt_max * 128;
    p.idesc = umma_idesc(f32);
    p.row0 = 0; p.ntiles = P.ntiles; p.slot_base = 0; p.bisect_first = 0;
    p.range = 1; p.nq_dev = c.list_len; p.ovf = c.ovf;
    cudaError_t e = umma_kernel_launch(S, c, f32, P.grid_range, *tmq, *tmqlo, *tmx, p, st);
    if (e != cudaSuccess) { err = std::string("tcgen05 range launch failed: ") + cudaGetErrorString(e); return -2; }
    ++*launches;
  }
  return 0;
}

}  // namespace b2r
