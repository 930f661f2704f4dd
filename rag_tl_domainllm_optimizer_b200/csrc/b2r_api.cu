// b2r_api.cu — the C ABI (include/b2r.h) over the sm_100a kernels.  No torch types, no CPU
// fallback: every compute entry point needs a CUDA device and fails loudly without one.
#include <cuda_runtime.h>
#include <cuda.h>
#include <dlfcn.h>
#include <float.h>
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <mutex>
#include <new>
#include <string>

#include "../../include/b2r.h"
#include "common.cuh"
#include "finalize.cuh"
#include "prep.cuh"
#include "score_exact.cuh"
#include "score_fma.cuh"
#include "score_umma.cuh"

using namespace b2r;

// --------------------------------------------------------------------------------------
// errors
// --------------------------------------------------------------------------------------
static thread_local std::string g_err;

static int fail(int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  g_err = buf;
  return code;
}

#define CU_TRY(expr)                                                                              \
  do {                                                                                            \
    cudaError_t e_ = (expr);                                                                      \
    if (e_ != cudaSuccess)                                                                        \
      return fail(e_ == cudaErrorMemoryAllocation ? B2R_ERR_OOM : B2R_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, \
                  cudaGetErrorString(e_), __FILE__, __LINE__);                                    \
  } while (0)

#define RC_TRY(expr)            \
  do {                          \
    int rc_ = (expr);           \
    if (rc_ != B2R_OK) return rc_; \
  } while (0)

extern "C" const char* b2r_last_error(void) { return g_err.c_str(); }
extern "C" const char* b2r_version(void) { return "b2r 0.1 sm_100a"; }
extern "C" int b2r_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}

// --------------------------------------------------------------------------------------
// small helpers
// --------------------------------------------------------------------------------------
static bool is_device_ptr(const void* p, int* owner = nullptr) {
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  if (owner) *owner = a.device;
  return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

struct DevBuf {
  void* p = nullptr;
  size_t bytes = 0;
  int ensure(size_t need) {
    if (need <= bytes) return B2R_OK;
    if (p) cudaFree(p);
    p = nullptr;
    bytes = 0;
    size_t want = need + need / 4;
    cudaError_t e = cudaMalloc(&p, want);
    if (e != cudaSuccess) {
      cudaGetLastError();
      e = cudaMalloc(&p, need);
      want = need;
    }
    if (e != cudaSuccess) {
      cudaGetLastError();
      p = nullptr;
      return fail(B2R_ERR_OOM, "device allocation of %zu bytes failed", need);
    }
    bytes = want;
    return B2R_OK;
  }
  void release() {
    if (p) cudaFree(p);
    p = nullptr;
    bytes = 0;
  }
};

static int round_up(int v, int m) { return (v + m - 1) / m * m; }
static int host_next_pow2(int v) {
  int p = 1;
  while (p < v) p <<= 1;
  return p;
}

struct DeviceGuard {
  int prev = -1;
  bool ok = false;
  explicit DeviceGuard(int dev) {
    if (cudaGetDevice(&prev) != cudaSuccess) { cudaGetLastError(); prev = -1; }
    ok = cudaSetDevice(dev) == cudaSuccess;
    if (!ok) cudaGetLastError();
  }
  ~DeviceGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

// --------------------------------------------------------------------------------------
// the handle
// --------------------------------------------------------------------------------------
struct NcclApi;   // dlopen'ed entry points of the NCCL library (sharded search), below

struct b2r_index {
  int64_t d = 0;
  int metric = 0, storage = 0, device = 0;
  int ld = 0;             // row pitch in elements (multiple of 64: one 128-byte TMA/UMMA swizzle row in bf16, two in fp32)
  int64_t ntotal = 0, cap_rows = 0;
  void* X = nullptr;      // [cap_rows][ld] fp32 or bf16
  void* Xs = nullptr;     // optional bf16 shadow of fp32 rows, same pitch (b2r_index_set_shadow)
  bool shadow = false;
  float* norm2 = nullptr; // [cap_rows] squared norm of the stored rows (L2 metric only)
  float* maxnorm2 = nullptr;  // device scalar: largest stored squared norm
  int sm_count = 148;
  int engine = B2R_ENGINE_AUTO, pad = 6;
  bool verify = true;                  // unproven queries go on to the range pass and the exhaustive scan (device side)
  // per-search workspace
  DevBuf qf, qh, qlo, qn2, thr, cand, cnt, stats, stage_q, stage_d, stage_i, stage_x, misc;
  DevBuf fb;                           // hand-over lists of unproven queries: 16 counters, then 4 arrays of nq words
  DevBuf qc_f, qc_lo, qc_h;            // their prepared rows, compacted, as the range pass reads them
  DevBuf qc2_f, qc2_lo;                // ... and as the second (3xTF32) range pass reads them
  DevBuf exl;                          // per-CTA top-k lists of the exhaustive scan
  DevBuf shard_local, shard_all;       // sharded search: this rank's packed top-k, and every rank's
  static constexpr int kEvPool = 256;           // scoring-kernel timers: one event pair per recent search
  cudaEvent_t ev0[kEvPool] = {}, ev1[kEvPool] = {};
  cudaEvent_t ev_done = nullptr;                // recorded after the last kernel of the most recent search
  cudaStream_t side = nullptr;                  // non-blocking stream for the counters' read-back
  int64_t nrec = 0;                             // searches timed since the last timer reset
  bool have_times = false, searched = false;
  int exact_occ[2] = {0, 0};                    // resident CTAs per SM of score_exact_kernel (per storage), 0 = unknown
  size_t exact_occ_smem[2] = {0, 0};
  b2r_search_stats last{};
  UmmaState umma;         // tensor maps and scratch of the tcgen05 engine
  // row-sharded search (b2r_index_set_comm / b2r_index_comm_init)
  void* comm = nullptr;
  bool comm_owned = false;
  int rank = 0, world = 1;
  int64_t row_offset = 0;
  std::mutex mu;
  size_t elem_bytes() const { return storage == B2R_STORE_BF16 ? 2 : 4; }
};

extern "C" int b2r_index_create(b2r_index** out, int64_t d, int metric, int storage, int device) {
  if (!out) return fail(B2R_ERR_ARG, "out is null");
  *out = nullptr;
  if (d <= 0 || d > 65536) return fail(B2R_ERR_ARG, "dimension %lld out of range (1..65536)", (long long)d);
  if (metric != B2R_METRIC_IP && metric != B2R_METRIC_L2) return fail(B2R_ERR_ARG, "unknown metric %d", metric);
  if (storage != B2R_STORE_F32 && storage != B2R_STORE_BF16) return fail(B2R_ERR_ARG, "unknown storage %d", storage);
  int ndev = b2r_device_count();
  if (ndev <= 0) return fail(B2R_ERR_CUDA, "no CUDA device visible: this library has no CPU path");
  if (device < 0 || device >= ndev) return fail(B2R_ERR_ARG, "device %d out of range (0..%d)", device, ndev - 1);
  DeviceGuard g(device);
  if (!g.ok) return fail(B2R_ERR_CUDA, "cudaSetDevice(%d) failed", device);
  cudaDeviceProp prop;
  CU_TRY(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10) return fail(B2R_ERR_CUDA, "device %d is sm_%d%d; this library is built for sm_100a only", device, prop.major, prop.minor);
  b2r_index* idx = new (std::nothrow) b2r_index();
  if (!idx) return fail(B2R_ERR_OOM, "host allocation failed");
  idx->d = d;
  idx->metric = metric;
  idx->storage = storage;
  idx->device = device;
  idx->ld = round_up(int(d), 64);
  idx->sm_count = prop.multiProcessorCount;
  cudaError_t e = cudaMalloc(&idx->maxnorm2, sizeof(float));
  if (e == cudaSuccess) e = cudaMemset(idx->maxnorm2, 0, sizeof(float));
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&idx->ev_done, cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&idx->side, cudaStreamNonBlocking);
  for (int i = 0; i < b2r_index::kEvPool && e == cudaSuccess; ++i) {
    e = cudaEventCreate(&idx->ev0[i]);
    if (e == cudaSuccess) e = cudaEventCreate(&idx->ev1[i]);
  }
  if (e != cudaSuccess) {
    cudaGetLastError();
    b2r_index_free(idx);
    return fail(B2R_ERR_CUDA, "handle set-up failed: %s", cudaGetErrorString(e));
  }
  *out = idx;
  return B2R_OK;
}

static void comm_release(b2r_index* idx);

extern "C" void b2r_index_free(b2r_index* idx) {
  if (!idx) return;
  DeviceGuard g(idx->device);
  cudaDeviceSynchronize();
  comm_release(idx);
  if (idx->X) cudaFree(idx->X);
  if (idx->Xs) cudaFree(idx->Xs);
  if (idx->norm2) cudaFree(idx->norm2);
  if (idx->maxnorm2) cudaFree(idx->maxnorm2);
  DevBuf* bufs[] = {&idx->qf, &idx->qh, &idx->qlo, &idx->qn2, &idx->thr, &idx->cand, &idx->cnt, &idx->stats,
                    &idx->stage_q, &idx->stage_d, &idx->stage_i, &idx->stage_x, &idx->misc, &idx->fb, &idx->qc_f,
                    &idx->qc_lo, &idx->qc_h, &idx->qc2_f, &idx->qc2_lo, &idx->exl, &idx->shard_local, &idx->shard_all};
  for (DevBuf* b : bufs) b->release();
  idx->umma.release();
  for (int i = 0; i < b2r_index::kEvPool; ++i) {
    if (idx->ev0[i]) cudaEventDestroy(idx->ev0[i]);
    if (idx->ev1[i]) cudaEventDestroy(idx->ev1[i]);
  }
  if (idx->ev_done) cudaEventDestroy(idx->ev_done);
  if (idx->side) cudaStreamDestroy(idx->side);
  cudaGetLastError();
  delete idx;
}

static int grow(b2r_index* idx, int64_t rows, cudaStream_t st) {
  if (rows <= idx->cap_rows) return B2R_OK;
  int64_t want = rows;
  if (idx->cap_rows > 0) want = std::max<int64_t>(rows, idx->cap_rows + idx->cap_rows / 2);
  const size_t rowb = size_t(idx->ld) * idx->elem_bytes();
  void* nx = nullptr;
  float* nn = nullptr;
  void* ns = nullptr;
  auto release_new = [&]() {
    if (nx) cudaFree(nx);
    if (nn) cudaFree(nn);
    if (ns) cudaFree(ns);
    cudaGetLastError();
  };
  cudaError_t e = cudaMalloc(&nx, size_t(want) * rowb);
  if (e != cudaSuccess && want > rows) {
    cudaGetLastError();
    want = rows;
    e = cudaMalloc(&nx, size_t(want) * rowb);
  }
  if (e != cudaSuccess) {
    cudaGetLastError();
    return fail(B2R_ERR_OOM, "cannot allocate %zu bytes for %lld rows", size_t(want) * rowb, (long long)want);
  }
  if (idx->metric == B2R_METRIC_L2 && cudaMalloc(&nn, size_t(want) * sizeof(float)) != cudaSuccess) {
    release_new();
    return fail(B2R_ERR_OOM, "cannot allocate row norms for %lld rows", (long long)want);
  }
  if (idx->shadow && cudaMalloc(&ns, size_t(want) * idx->ld * 2) != cudaSuccess) {
    release_new();
    return fail(B2R_ERR_OOM, "cannot allocate the bf16 shadow for %lld rows", (long long)want);
  }
  if (idx->ntotal > 0) {
    // rows added earlier may still be in flight on another stream of the caller: growth is rare, wait for the device
    e = cudaDeviceSynchronize();
    if (e == cudaSuccess) e = cudaMemcpyAsync(nx, idx->X, size_t(idx->ntotal) * rowb, cudaMemcpyDeviceToDevice, st);
    if (e == cudaSuccess && nn) e = cudaMemcpyAsync(nn, idx->norm2, size_t(idx->ntotal) * sizeof(float), cudaMemcpyDeviceToDevice, st);
    if (e == cudaSuccess && ns) e = cudaMemcpyAsync(ns, idx->Xs, size_t(idx->ntotal) * idx->ld * 2, cudaMemcpyDeviceToDevice, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) {
      release_new();
      return fail(B2R_ERR_CUDA, "copying %lld rows into the grown storage failed: %s", (long long)idx->ntotal, cudaGetErrorString(e));
    }
  }
  if (idx->X) cudaFree(idx->X);
  if (idx->Xs) cudaFree(idx->Xs);
  if (idx->norm2) cudaFree(idx->norm2);
  idx->X = nx;
  idx->Xs = ns;
  idx->norm2 = nn;
  idx->cap_rows = want;
  idx->umma.invalidate();
  return B2R_OK;
}

extern "C" int b2r_index_reserve(b2r_index* idx, int64_t n) {
  if (!idx) return fail(B2R_ERR_ARG, "index is null");
  if (n < 0) return fail(B2R_ERR_ARG, "negative row count");
  std::lock_guard<std::mutex> lk(idx->mu);
  DeviceGuard g(idx->device);
  return grow(idx, n, nullptr);
}

extern "C" int b2r_index_add(b2r_index* idx, int64_t n, const float* x, int normalize, void* stream) {
  if (!idx) return fail(B2R_ERR_ARG, "index is null");
  if (n < 0) return fail(B2R_ERR_ARG, "negative row count");
  if (n == 0) return B2R_OK;
  if (!x) return fail(B2R_ERR_ARG, "x is null");
  std::lock_guard<std::mutex> lk(idx->mu);
  if (idx->ntotal + n > 0xfffffffeLL) return fail(B2R_ERR_LIMIT, "a shard holds at most 2^32-2 rows");
  DeviceGuard g(idx->device);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  RC_TRY(grow(idx, idx->ntotal + n, st));
  const bool dev_in = is_device_ptr(x);
  const int64_t chunk = dev_in ? n : std::max<int64_t>(1, std::min<int64_t>(n, (int64_t(256) << 20) / (idx->d * 4)));
  if (!dev_in) RC_TRY(idx->stage_x.ensure(size_t(chunk) * idx->d * 4));
  for (int64_t r0 = 0; r0 < n; r0 += chunk) {
    const int64_t m = std::min(chunk, n - r0);
    const float* src = x + r0 * idx->d;
    if (!dev_in) {
      CU_TRY(cudaMemcpyAsync(idx->stage_x.p, src, size_t(m) * idx->d * 4, cudaMemcpyHostToDevice, st));
      src = static_cast<const float*>(idx->stage_x.p);
    }
    const int64_t row0 = idx->ntotal + r0;
    const int blocks = int(std::min<int64_t>((m + 7) / 8, int64_t(idx->sm_count) * 16));
    float* n2 = idx->norm2 ? idx->norm2 + row0 : nullptr;
    if (idx->storage == B2R_STORE_BF16) {
      __nv_bfloat16* dst = static_cast<__nv_bfloat16*>(idx->X) + row0 * idx->ld;
      ingest_rows_kernel<__nv_bfloat16><<<blocks, 256, 0, st>>>(src, m, int(idx->d), dst, idx->ld, normalize, n2, idx->maxnorm2);
    } else {
      float* dst = static_cast<float*>(idx->X) + row0 * idx->ld;
      ingest_rows_kernel<float><<<blocks, 256, 0, st>>>(src, m, int(idx->d), dst, idx->ld, normalize, n2, idx->maxnorm2);
      if (idx->shadow)
        shadow_rows_kernel<<<blocks, 256, 0, st>>>(dst, m, idx->ld, static_cast<__nv_bfloat16*>(idx->Xs) + row0 * idx->ld);
    }
    CU_TRY(cudaGetLastError());
    if (!dev_in) CU_TRY(cudaStreamSynchronize(st));  // the staging buffer is reused by the next chunk
  }
  idx->ntotal += n;
  return B2R_OK;
}

extern "C" int b2r_index_reset(b2r_index* idx) {
  if (!idx) return fail(B2R_ERR_ARG, "index is null");
  std::lock_guard<std::mutex> lk(idx->mu);
  DeviceGuard g(idx->device);
  CU_TRY(cudaDeviceSynchronize());
  idx->ntotal = 0;
  CU_TRY(cudaMemset(idx->maxnorm2, 0, sizeof(float)));
  return B2R_OK;
}

extern "C" int b2r_index_ntotal(const b2r_index* idx, int64_t* out) {
  if (!idx || !out) return fail(B2R_ERR_ARG, "null argument");
  std::lock_guard<std::mutex> lk(const_cast<b2r_index*>(idx)->mu);
  *out = idx->ntotal;
  return B2R_OK;
}

template <typename T>
__global__ void reconstruct_kernel(const T* row, int d, float* out) {
  for (int j = threadIdx.x; j < d; j += blockDim.x) out[j] = ld_elem(row, j);
}

extern "C" int b2r_index_reconstruct(b2r_index* idx, int64_t i, float* out, void* stream) {
  if (!idx || !out) return fail(B2R_ERR_ARG, "null argument");
  std::lock_guard<std::mutex> lk(idx->mu);
  if (i < 0 || i >= idx->ntotal) return fail(B2R_ERR_ARG, "row %lld out of range (ntotal %lld)", (long long)i, (long long)idx->ntotal);
  DeviceGuard g(idx->device);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const bool dev_out = is_device_ptr(out);
  float* dst = out;
  if (!dev_out) {
    RC_TRY(idx->misc.ensure(size_t(idx->d) * 4));
    dst = static_cast<float*>(idx->misc.p);
  }
  if (idx->storage == B2R_STORE_BF16)
    reconstruct_kernel<<<1, 256, 0, st>>>(static_cast<const __nv_bfloat16*>(idx->X) + i * idx->ld, int(idx->d), dst);
  else
    reconstruct_kernel<<<1, 256, 0, st>>>(static_cast<const float*>(idx->X) + i * idx->ld, int(idx->d), dst);
  CU_TRY(cudaGetLastError());
  if (!dev_out) {
    CU_TRY(cudaMemcpyAsync(out, dst, size_t(idx->d) * 4, cudaMemcpyDeviceToHost, st));
    CU_TRY(cudaStreamSynchronize(st));
  }
  return B2R_OK;
}

extern "C" int b2r_index_set_engine(b2r_index* idx, int engine) {
  if (!idx) return fail(B2R_ERR_ARG, "index is null");
  if (engine < B2R_ENGINE_AUTO || engine > B2R_ENGINE_UMMA3) return fail(B2R_ERR_ARG, "unknown engine %d", engine);
  std::lock_guard<std::mutex> lk(idx->mu);
  idx->engine = engine;
  return B2R_OK;
}

extern "C" int b2r_index_set_shadow(b2r_index* idx, int on) {
  if (!idx) return fail(B2R_ERR_ARG, "index is null");
  if (idx->storage != B2R_STORE_F32) return fail(B2R_ERR_ARG, "a bf16 shadow only makes sense for fp32 rows (exact mode)");
  std::lock_guard<std::mutex> lk(idx->mu);
  if (idx->ntotal != 0 || idx->cap_rows != 0) return fail(B2R_ERR_STATE, "set the shadow option before the first add/reserve");
  idx->shadow = on != 0;
  return B2R_OK;
}

extern "C" int b2r_index_set_verify(b2r_index* idx, int on) {
  if (!idx) return fail(B2R_ERR_ARG, "index is null");
  std::lock_guard<std::mutex> lk(idx->mu);
  idx->verify = on != 0;
  return B2R_OK;
}

extern "C" int b2r_index_set_pad(b2r_index* idx, int pad) {
  if (!idx) return fail(B2R_ERR_ARG, "index is null");
  if (pad < 0 || pad > 512) return fail(B2R_ERR_ARG, "pad %d out of range (0..512)", pad);
  std::lock_guard<std::mutex> lk(idx->mu);
  idx->pad = pad;
  return B2R_OK;
}

// --------------------------------------------------------------------------------------
// search
// --------------------------------------------------------------------------------------
static int launch_fma(b2r_index* idx, int64_t nq, int kp, cudaStream_t st, FinalizeParams& fp, int* launches) {
  const int64_t ntiles = (idx->ntotal + kFmaTile - 1) / kFmaTile;
  const int grid = int(std::max<int64_t>(1, std::min<int64_t>(ntiles, int64_t(idx->sm_count) * 2)));
  int cap = host_next_pow2(std::max(2 * kp, kp + 2 * kFmaTile));
  size_t smem = size_t(kFmaQB) * idx->ld * 4 + size_t(kFmaQB) * cap * 8;
  if (smem > 226 * 1024) {   // large k: settle for the smallest buffer that still absorbs one tile between compactions
    cap = host_next_pow2(kp + 2 * kFmaTile);
    smem = size_t(kFmaQB) * idx->ld * 4 + size_t(kFmaQB) * cap * 8;
  }
  if (smem > 226 * 1024)
    return fail(B2R_ERR_LIMIT, "FMA engine needs %zu bytes of shared memory (d=%lld, k'=%d); limit is 232448", smem, (long long)idx->d, kp);
  RC_TRY(idx->cand.ensure(size_t(grid) * nq * kp * 8));
  FmaParams p{};
  p.X = idx->X; p.n = idx->ntotal; p.ld = idx->ld; p.d = int(idx->d);
  p.Q = static_cast<const float*>(idx->qf.p); p.ldq = idx->ld;
  p.norm2 = idx->metric == B2R_METRIC_L2 ? idx->norm2 : nullptr;
  p.thr = static_cast<uint32_t*>(idx->thr.p);
  p.cand = static_cast<uint64_t*>(idx->cand.p);
  p.nq = int(nq); p.kp = kp; p.cap = cap;
  {   // opt in to the full 227 KiB once per device and instance (never lowered: other handles may be launching)
    static std::mutex mu;
    static bool done[64][2] = {};
    std::lock_guard<std::mutex> lk(mu);
    const int which = idx->storage == B2R_STORE_BF16 ? 1 : 0;
    if (idx->device < 64 && !done[idx->device][which]) {
      if (which)
        CU_TRY(cudaFuncSetAttribute(score_fma_kernel<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
      else
        CU_TRY(cudaFuncSetAttribute(score_fma_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
      done[idx->device][which] = true;
    }
  }
  for (int64_t q0 = 0; q0 < nq; q0 += kFmaQB) {
    p.q0 = int(q0);
    p.nqb = int(std::min<int64_t>(kFmaQB, nq - q0));
    if (idx->storage == B2R_STORE_BF16)
      score_fma_kernel<__nv_bfloat16><<<grid, kFmaWarps * 32, smem, st>>>(p);
    else
      score_fma_kernel<float><<<grid, kFmaWarps * 32, smem, st>>>(p);
    ++*launches;
  }
  CU_TRY(cudaGetLastError());
  fp.cand = p.cand;
  fp.slot_stride = int64_t(nq) * kp;
  fp.q_stride = kp;
  fp.nslots = grid;
  fp.per = kp;
  fp.cnt = nullptr;
  fp.errfac = float((double(idx->d) + 8.0) * 6.1e-8);   // (d+8) * 2^-24 * 1.02
  return B2R_OK;
}

// Kernels whose dynamic shared memory can exceed the 48 KiB default need the opt-in attribute; it is set to a
// fixed size once per device and kernel instance, never lowered (other handles may be launching).
enum { kAttrFinalize = 0, kAttrRange = 2, kAttrExact = 4, kAttrMerge = 6, kAttrCount = 7 };
constexpr size_t kFinalizeSmemMax = 64 * 1024, kExactSmemMax = 160 * 1024;
static int smem_attr_once(int device, int which) {
  static std::mutex mu;
  static bool done[64][kAttrCount] = {};
  std::lock_guard<std::mutex> lk(mu);
  if (device < 0 || device >= 64) return fail(B2R_ERR_ARG, "device %d out of range", device);
  if (done[device][which]) return B2R_OK;
  switch (which) {
    case kAttrFinalize + 0: CU_TRY(cudaFuncSetAttribute(finalize_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(kFinalizeSmemMax))); break;
    case kAttrFinalize + 1: CU_TRY(cudaFuncSetAttribute(finalize_kernel<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(kFinalizeSmemMax))); break;
    case kAttrRange + 0: CU_TRY(cudaFuncSetAttribute(finalize_range_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(kFinalizeSmemMax))); break;
    case kAttrRange + 1: CU_TRY(cudaFuncSetAttribute(finalize_range_kernel<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(kFinalizeSmemMax))); break;
    case kAttrExact + 0: CU_TRY(cudaFuncSetAttribute(score_exact_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(kExactSmemMax))); break;
    case kAttrExact + 1: CU_TRY(cudaFuncSetAttribute(score_exact_kernel<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(kExactSmemMax))); break;
    case kAttrMerge: CU_TRY(cudaFuncSetAttribute(merge_packed_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(kFinalizeSmemMax))); break;
    default: return fail(B2R_ERR_ARG, "unknown kernel %d", which);
  }
  done[device][which] = true;
  return B2R_OK;
}

constexpr int kRangeKeep = 2048;   // most candidates finalize_range_kernel re-scores per query
constexpr int kFbHeader = 16;      // counters in front of the hand-over lists

// One batch (<= 8192 queries), device pointers only, everything enqueued on `st`, nothing read back:
//   prep -> first pass (tcgen05 seed/select/main, or FMA) -> finalize (+ proof, hand-over lists)
//        -> [range pass -> finalize_range] -> exhaustive scan of whatever is still unproven.
// The last three launches return at once when their list is empty.
static int search_device(b2r_index* idx, int64_t nq, const float* q_dev, int64_t k, int normalize_queries,
                         int64_t row_offset, float* D, int64_t* I, uint64_t* packed, cudaStream_t st, bool clear_stats) {
  // engine and K'.  The 1xTF32 tensor-core pass truncates both operands to 10 mantissa bits, so it keeps a wider
  // candidate list (K' >= 2k, at least k + 22) for the completeness proof to go through; FMA, bf16 (exact
  // products) and 3xTF32 keep k + pad.  The pad grows with k: gaps between neighbouring scores shrink like 1/rank.
  const int engine = idx->engine;
  bool use_umma = false, split = false, use_shadow = false;
  int kp = int(std::min<int64_t>(k + std::max<int64_t>(idx->pad, k / 8), 2048));
  if (idx->ntotal > 0 && engine != B2R_ENGINE_FMA) {
    split = engine == B2R_ENGINE_UMMA3;
    if (split && idx->storage != B2R_STORE_F32)
      return fail(B2R_ERR_ARG, "the 3xTF32 engine needs fp32 rows (exact mode)");
    // exact mode with a bf16 shadow: score the shadow (half the bytes, twice the MMA rate); rounding both operands
    // to bf16 costs 2^-8 relative, twice the TF32 truncation, hence a slightly wider K'
    use_shadow = idx->shadow && idx->storage == B2R_STORE_F32 && !split && engine == B2R_ENGINE_AUTO;
    const int kp_tc = (idx->storage == B2R_STORE_F32 && !split)
                          ? int(k + std::max<int64_t>(std::max(idx->pad, use_shadow ? 30 : 22), k)) : kp;
    use_umma = umma_supported(idx->storage, int(idx->d), kp_tc, idx->ntotal);
    if (use_umma) kp = kp_tc;
    else if (engine == B2R_ENGINE_UMMA || engine == B2R_ENGINE_UMMA3)
      return fail(B2R_ERR_LIMIT, "the tcgen05 engine does not take this shape (storage %d, d %lld, k' %d)", idx->storage, (long long)idx->d, kp_tc);
    if (!use_umma) { split = false; use_shadow = false; }
  }
  const bool f32 = idx->storage == B2R_STORE_F32;
  // unproven queries are handed on (device side) when the proof can fail at all: rows may only be filtered out
  // when there are more of them than K'.  Range passes: fp32 rows searched with one TF32 pass (or its bf16 shadow)
  // get a 1xTF32 range pass (error band 2e-3: cheap, HBM-bound for up to 128 queries) and, for what does not fit,
  // a 3xTF32 one (band 7e-5); a forced 3xTF32 engine and bf16 rows get one range pass with their own engine.
  const bool chain = idx->verify && idx->ntotal > kp;
  const bool range_ok = chain && use_umma;
  const bool range2 = range_ok && f32 && !split;
  int launches = 0;
  const int ldq = idx->ld;
  RC_TRY(idx->qf.ensure(size_t(nq + 256) * ldq * 4));
  RC_TRY(idx->qh.ensure(size_t(nq + 256) * ldq * 2));
  if (split) RC_TRY(idx->qlo.ensure(size_t(nq + 256) * ldq * 4));
  RC_TRY(idx->qn2.ensure(size_t(nq) * 4));
  RC_TRY(idx->thr.ensure(size_t(nq) * 4));
  RC_TRY(idx->stats.ensure(4 * sizeof(unsigned long long)));
  RC_TRY(idx->fb.ensure(size_t(kFbHeader + 9 * nq) * 4));
  if (range_ok) {
    if (f32) {
      RC_TRY(idx->qc_f.ensure(size_t(nq + 256) * ldq * 4));
      if (split) RC_TRY(idx->qc_lo.ensure(size_t(nq + 256) * ldq * 4));
    } else {
      RC_TRY(idx->qc_h.ensure(size_t(nq + 256) * ldq * 2));
    }
    if (range2) {
      RC_TRY(idx->qc2_f.ensure(size_t(nq + 256) * ldq * 4));
      RC_TRY(idx->qc2_lo.ensure(size_t(nq + 256) * ldq * 4));
    }
  }
  uint32_t* fb = static_cast<uint32_t*>(idx->fb.p);
  {
    const int blocks = int(std::min<int64_t>((nq + 7) / 8, int64_t(idx->sm_count) * 8));
    CU_TRY(launch_pdl(prep_queries_kernel, blocks, 256, 0, st, q_dev, nq, int(idx->d), ldq, normalize_queries,
                      int(idx->storage == B2R_STORE_BF16), static_cast<float*>(idx->qf.p),
                      (idx->storage == B2R_STORE_BF16 || use_shadow) ? static_cast<__nv_bfloat16*>(idx->qh.p) : nullptr,
                      static_cast<float*>(idx->qn2.p), split ? static_cast<float*>(idx->qlo.p) : nullptr,
                      static_cast<uint32_t*>(idx->thr.p),
                      clear_stats ? static_cast<unsigned long long*>(idx->stats.p) : nullptr, fb));
    CU_TRY(cudaGetLastError());
    ++launches;
  }
  FinalizeParams fp{};
  int engine_used = B2R_ENGINE_FMA;
  int64_t main_rows = idx->ntotal;
  UmmaCall uc{};
  if (idx->ntotal > 0) {
    const int slot = int(idx->nrec % b2r_index::kEvPool);
    if (!use_umma) CU_TRY(cudaEventRecord(idx->ev0[slot], st));
    if (use_umma) {
      uc.X = use_shadow ? idx->Xs : idx->X; uc.n = idx->ntotal; uc.ld = idx->ld; uc.d = int(idx->d);
      uc.storage = use_shadow ? B2R_STORE_BF16 : idx->storage;
      uc.shadow = use_shadow;
      uc.Qh = idx->qh.p; uc.Qf = static_cast<const float*>(idx->qf.p); uc.nq = int(nq); uc.kp = kp;
      uc.norm2 = idx->metric == B2R_METRIC_L2 ? idx->norm2 : nullptr;
      uc.thr = static_cast<uint32_t*>(idx->thr.p);
      uc.sm_count = idx->sm_count;
      uc.split = split; uc.Qlo = split ? static_cast<const float*>(idx->qlo.p) : nullptr;
      uc.ev_main0 = idx->ev0[slot]; uc.ev_main1 = idx->ev1[slot];   // the timers bracket the main launch
      std::string err;
      int rc = umma_launch(idx->umma, uc, st, fp, &launches, err);
      if (rc != B2R_OK) return fail(rc, "%s", err.c_str());
      engine_used = split ? B2R_ENGINE_UMMA3 : use_shadow ? B2R_ENGINE_SHADOW : B2R_ENGINE_UMMA;
      main_rows = uc.main_rows;
      if (main_rows == 0) {   // tiny index: everything ran in the seed launch; nothing to time separately
        CU_TRY(cudaEventRecord(idx->ev0[slot], st));
        CU_TRY(cudaEventRecord(idx->ev1[slot], st));
      }
    } else {
      RC_TRY(launch_fma(idx, nq, kp, st, fp, &launches));
      CU_TRY(cudaEventRecord(idx->ev1[slot], st));
    }
    ++idx->nrec;
    idx->have_times = true;
  } else {
    idx->have_times = false;
    fp.nslots = 0;
    fp.per = 1;
  }
  fp.thr = static_cast<const uint32_t*>(idx->thr.p);
  fp.nq = int(nq); fp.kp = kp; fp.k = int(k);
  // few queries: one big CTA per query (latency: more warps re-scoring the K' rows at once); many queries: small CTAs,
  // several per SM (throughput)
  const int fin_threads = nq <= idx->sm_count ? kFinThreads : nq <= 4 * idx->sm_count ? 512 : 256;
  fp.sortn = std::max(4096, host_next_pow2(kp + 2048));   // gather rounds append up to 2048 keys
  fp.X = idx->X; fp.ld = idx->ld; fp.d = int(idx->d); fp.storage = idx->storage;
  fp.Q = static_cast<const float*>(idx->qf.p); fp.ldq = ldq; fp.qn2 = static_cast<const float*>(idx->qn2.p);
  fp.metric = idx->metric; fp.n = idx->ntotal; fp.row_offset = row_offset;
  fp.maxnorm2 = idx->maxnorm2;
  fp.D = D; fp.I = I; fp.packed = packed;
  fp.stats = static_cast<unsigned long long*>(idx->stats.p);
  FbList list1{}, list2{};
  if (chain) {
    fp.ex_n = fb + 1; fp.ex_bar = fb + 2; fp.ex_q = fb + kFbHeader + 8 * nq;
    fp.range_keep = kRangeKeep;
    if (range_ok) {
      list1.n = fb + 0; list1.q = fb + kFbHeader; list1.thr = fb + kFbHeader + nq; list1.thr_next = fb + kFbHeader + 2 * nq;
      list1.ovf = fb + kFbHeader + 3 * nq;
      list1.qc_f = f32 ? static_cast<float*>(idx->qc_f.p) : nullptr;
      list1.qc_lo = (f32 && split) ? static_cast<float*>(idx->qc_lo.p) : nullptr;
      list1.qc_h = f32 ? nullptr : static_cast<__nv_bfloat16*>(idx->qc_h.p);
      fp.errfac_range = umma_errfac(f32, split, false, int(idx->d));
      fp.errfac_range2 = range2 ? umma_errfac(true, true, false, int(idx->d)) : fp.errfac_range;
      fp.next = list1;
    }
    if (range2) {
      list2.n = fb + 3; list2.q = fb + kFbHeader + 4 * nq; list2.thr = fb + kFbHeader + 5 * nq; list2.thr_next = fb + kFbHeader + 6 * nq;
      list2.ovf = fb + kFbHeader + 7 * nq;
      list2.qc_f = static_cast<float*>(idx->qc2_f.p);
      list2.qc_lo = static_cast<float*>(idx->qc2_lo.p);
    }
  }
  const size_t smem = size_t(fp.sortn + host_next_pow2(kp)) * 8;
  RC_TRY(smem_attr_once(idx->device, kAttrFinalize + (f32 ? 0 : 1)));
  if (f32) CU_TRY(launch_pdl(finalize_kernel<float>, int(nq), fin_threads, smem, st, fp));
  else CU_TRY(launch_pdl(finalize_kernel<__nv_bfloat16>, int(nq), fin_threads, smem, st, fp));
  CU_TRY(cudaGetLastError());
  ++launches;

  if (chain) {
    // one range pass: scoring launch(es) over the list's compact queries, then its merge
    auto range_pass = [&](const FbList& cur, const FbList& next, bool pass_split, int pass) -> int {
      UmmaCall rc_call{};
      rc_call.X = idx->X; rc_call.n = idx->ntotal; rc_call.ld = idx->ld; rc_call.d = int(idx->d);
      rc_call.storage = idx->storage;
      rc_call.Qh = cur.qc_h; rc_call.Qf = cur.qc_f; rc_call.Qlo = cur.qc_lo;
      rc_call.nq = int(nq); rc_call.kp = kp;
      rc_call.norm2 = uc.norm2; rc_call.thr = cur.thr; rc_call.sm_count = idx->sm_count;
      rc_call.split = pass_split;
      rc_call.list_len = cur.n; rc_call.ovf = cur.ovf;
      FinalizeParams rp = fp;
      rp.cur = cur; rp.next = next;
      std::string err;
      int rc = umma_launch_range(idx->umma, rc_call, uc, pass, st, rp, &launches, err);
      if (rc != B2R_OK) return fail(rc, "%s", err.c_str());
      rp.sortn = 2 * kRangeKeep;
      const size_t rsmem = size_t(rp.sortn + kRangeKeep) * 8;
      RC_TRY(smem_attr_once(idx->device, kAttrRange + (f32 ? 0 : 1)));
      const int rgrid = int(std::min<int64_t>(nq, int64_t(idx->sm_count) * 2));
      if (f32) CU_TRY(launch_pdl(finalize_range_kernel<float>, rgrid, kFinThreads, rsmem, st, rp));
      else CU_TRY(launch_pdl(finalize_range_kernel<__nv_bfloat16>, rgrid, kFinThreads, rsmem, st, rp));
      CU_TRY(cudaGetLastError());
      ++launches;
      return B2R_OK;
    };
    if (range_ok) RC_TRY(range_pass(list1, range2 ? list2 : FbList{}, split, 1));
    if (range2) RC_TRY(range_pass(list2, FbList{}, true, 2));
    // exhaustive DOT64 scan of whatever is still unproven (cooperative: its CTAs meet at a grid barrier)
    FinalizeParams xp = fp;
    xp.sortn = std::max(4096, host_next_pow2(int(k) + 2048));
    int cap = 128;
    while (cap < int(k) + 64) cap <<= 1;
    const size_t xsmem = std::max(size_t(kExQB) * cap * 8, size_t(xp.sortn + host_next_pow2(int(k))) * 8);
    const int w = f32 ? 0 : 1;
    RC_TRY(smem_attr_once(idx->device, kAttrExact + w));
    if (idx->exact_occ[w] == 0 || idx->exact_occ_smem[w] != xsmem) {
      int occ = 0;
      if (f32) CU_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, score_exact_kernel<float>, kExThreads, xsmem));
      else CU_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, score_exact_kernel<__nv_bfloat16>, kExThreads, xsmem));
      if (occ < 1) return fail(B2R_ERR_LIMIT, "the exhaustive scan does not fit an SM (k=%lld)", (long long)k);
      idx->exact_occ[w] = std::min(occ, 2);
      idx->exact_occ_smem[w] = xsmem;
    }
    const int xgrid = idx->sm_count * idx->exact_occ[w];
    int64_t wave = (int64_t(32) << 20) / (int64_t(xgrid) * k * 8) / kExQB * kExQB;
    wave = std::max<int64_t>(kExQB, std::min<int64_t>(wave, (nq + kExQB - 1) / kExQB * kExQB));
    RC_TRY(idx->exl.ensure(size_t(wave) * xgrid * k * 8));
    uint64_t* lists = static_cast<uint64_t*>(idx->exl.p);
    int wave_i = int(wave);
    void* args[] = {&xp, &lists, &wave_i, &cap};
    if (f32) CU_TRY(cudaLaunchCooperativeKernel(reinterpret_cast<void*>(score_exact_kernel<float>), dim3(xgrid), dim3(kExThreads), args, xsmem, st));
    else CU_TRY(cudaLaunchCooperativeKernel(reinterpret_cast<void*>(score_exact_kernel<__nv_bfloat16>), dim3(xgrid), dim3(kExThreads), args, xsmem, st));
    ++launches;
  }
  if (clear_stats) {
    idx->last = b2r_search_stats{};
    idx->last.launches = 0;
  }
  idx->last.nq += nq; idx->last.k = k; idx->last.kprime = kp;
  idx->last.engine = engine_used; idx->last.launches += launches; idx->last.main_rows = main_rows;
  return B2R_OK;
}

static int search_batch(b2r_index* idx, int64_t nq, const float* q, int64_t k, int normalize_queries, int64_t row_offset,
                        float* D, int64_t* I, uint64_t* packed, cudaStream_t st, bool first_batch) {
  const bool q_dev = is_device_ptr(q);
  const bool out_dev = packed ? true : is_device_ptr(D);
  if (!packed && out_dev != is_device_ptr(I)) return fail(B2R_ERR_ARG, "D and I must both be host or both be device pointers");
  if (packed && !is_device_ptr(packed)) return fail(B2R_ERR_ARG, "packed must be a device pointer");
  const float* qd = q;
  if (!q_dev) {
    RC_TRY(idx->stage_q.ensure(size_t(nq) * idx->d * 4));
    CU_TRY(cudaMemcpyAsync(idx->stage_q.p, q, size_t(nq) * idx->d * 4, cudaMemcpyHostToDevice, st));
    qd = static_cast<const float*>(idx->stage_q.p);
  }
  float* Dd = D;
  int64_t* Id = I;
  const bool host_out = !packed && !out_dev;
  if (host_out) {
    RC_TRY(idx->stage_d.ensure(size_t(nq) * k * 4));
    RC_TRY(idx->stage_i.ensure(size_t(nq) * k * 8));
    Dd = static_cast<float*>(idx->stage_d.p);
    Id = static_cast<int64_t*>(idx->stage_i.p);
  }
  RC_TRY(search_device(idx, nq, qd, k, normalize_queries, row_offset, Dd, Id, packed, st, first_batch));
  if (host_out) {
    CU_TRY(cudaMemcpyAsync(D, Dd, size_t(nq) * k * 4, cudaMemcpyDeviceToHost, st));
    CU_TRY(cudaMemcpyAsync(I, Id, size_t(nq) * k * 8, cudaMemcpyDeviceToHost, st));
  }
  // host buffers: the results must be there (and the query buffer reusable) on return; device pointers: enqueue only
  if (host_out || !q_dev) CU_TRY(cudaStreamSynchronize(st));
  return B2R_OK;
}

static int check_search_args(b2r_index* idx, int64_t nq, const float* q, int64_t k) {
  if (!idx) return fail(B2R_ERR_ARG, "index is null");
  if (nq < 0) return fail(B2R_ERR_ARG, "negative query count");
  if (k <= 0) return fail(B2R_ERR_ARG, "k must be positive (got %lld)", (long long)k);
  if (k > 1024) return fail(B2R_ERR_LIMIT, "k=%lld exceeds the supported maximum of 1024", (long long)k);
  if (nq > (int64_t(1) << 22)) return fail(B2R_ERR_LIMIT, "at most 2^22 queries per call");
  if (nq > 0 && !q) return fail(B2R_ERR_ARG, "q is null");
  return B2R_OK;
}

// Large query sets go through in batches of 8192 queries so that the candidate scratch stays bounded (the corpus is
// streamed once per 2048 queries either way).  The counters of the batches add up on the device.
constexpr int64_t kSearchBatch = 8192;
static int search_any(b2r_index* idx, int64_t nq, const float* q, int64_t k, int normalize_queries, int64_t row_offset,
                      float* D, int64_t* I, uint64_t* packed, void* stream) {
  RC_TRY(check_search_args(idx, nq, q, k));
  if (nq == 0) return B2R_OK;
  if (packed ? false : (!D || !I)) return fail(B2R_ERR_ARG, "D or I is null");
  std::lock_guard<std::mutex> lk(idx->mu);
  if (idx->ntotal + row_offset > 0xfffffffeLL || row_offset < 0) return fail(B2R_ERR_LIMIT, "global ids must stay below 2^32-1");
  DeviceGuard g(idx->device);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  for (int64_t q0 = 0; q0 < nq; q0 += kSearchBatch) {
    const int64_t m = std::min(kSearchBatch, nq - q0);
    RC_TRY(search_batch(idx, m, q + q0 * idx->d, k, normalize_queries, row_offset, D ? D + q0 * k : nullptr,
                        I ? I + q0 * k : nullptr, packed ? packed + q0 * k : nullptr, st, q0 == 0));
  }
  idx->searched = true;
  CU_TRY(cudaEventRecord(idx->ev_done, st));
  return B2R_OK;
}

extern "C" int b2r_index_search(b2r_index* idx, int64_t nq, const float* q, int64_t k, float* D, int64_t* I, void* stream) {
  return search_any(idx, nq, q, k, 0, 0, D, I, nullptr, stream);
}
extern "C" int b2r_index_search_ex(b2r_index* idx, int64_t nq, const float* q, int64_t k, int normalize_queries,
                                   int64_t row_offset, float* D, int64_t* I, void* stream) {
  return search_any(idx, nq, q, k, normalize_queries, row_offset, D, I, nullptr, stream);
}
extern "C" int b2r_index_search_packed(b2r_index* idx, int64_t nq, const float* q, int64_t k, int normalize_queries,
                                       int64_t row_offset, uint64_t* packed, void* stream) {
  if (!packed) return fail(B2R_ERR_ARG, "packed is null");
  return search_any(idx, nq, q, k, normalize_queries, row_offset, nullptr, nullptr, packed, stream);
}

static int merge_launch(int metric, int64_t G, int64_t nq, int64_t k, const uint64_t* packed, int64_t g_stride, float* D,
                        int64_t* I, int device, cudaStream_t st) {
  const int sortn = std::max(4096, host_next_pow2(int(k) + 2048));   // gather rounds append up to 2048 keys
  RC_TRY(smem_attr_once(device, kAttrMerge));
  CU_TRY(launch_pdl(merge_packed_kernel, int(nq), 256, size_t(sortn + host_next_pow2(int(k))) * 8, st, packed, int(G), int(nq),
                    int(k), g_stride, sortn, metric, D, I));
  CU_TRY(cudaGetLastError());
  return B2R_OK;
}

extern "C" int b2r_merge_packed_strided(int metric, int64_t G, int64_t nq, int64_t k, const uint64_t* packed,
                                        int64_t g_stride, float* D, int64_t* I, int device, void* stream) {
  if (g_stride < nq * k) return fail(B2R_ERR_ARG, "g_stride %lld is smaller than nq*k", (long long)g_stride);
  if (metric != B2R_METRIC_IP && metric != B2R_METRIC_L2) return fail(B2R_ERR_ARG, "unknown metric %d", metric);
  if (G <= 0 || nq < 0 || k <= 0) return fail(B2R_ERR_ARG, "bad shape G=%lld nq=%lld k=%lld", (long long)G, (long long)nq, (long long)k);
  if (k > 1024) return fail(B2R_ERR_LIMIT, "k=%lld exceeds the supported maximum of 1024", (long long)k);
  if (G > kFinMaxSlots) return fail(B2R_ERR_LIMIT, "at most %d lists can be merged (got %lld)", kFinMaxSlots, (long long)G);
  if (nq == 0) return B2R_OK;
  if (!packed || !D || !I) return fail(B2R_ERR_ARG, "null argument");
  if (b2r_device_count() <= 0) return fail(B2R_ERR_CUDA, "no CUDA device visible: this library has no CPU path");
  if (!is_device_ptr(packed) || !is_device_ptr(D) || !is_device_ptr(I)) return fail(B2R_ERR_ARG, "packed, D and I must be device pointers");
  DeviceGuard g(device);
  if (!g.ok) return fail(B2R_ERR_CUDA, "cudaSetDevice(%d) failed", device);
  return merge_launch(metric, G, nq, k, packed, g_stride, D, I, device, static_cast<cudaStream_t>(stream));
}

extern "C" int b2r_merge_packed(int metric, int64_t G, int64_t nq, int64_t k, const uint64_t* packed, float* D,
                                int64_t* I, int device, void* stream) {
  return b2r_merge_packed_strided(metric, G, nq, k, packed, nq * k, D, I, device, stream);
}

extern "C" int b2r_index_last_stats(b2r_index* idx, b2r_search_stats* out) {
  if (!idx || !out) return fail(B2R_ERR_ARG, "null argument");
  std::lock_guard<std::mutex> lk(idx->mu);
  DeviceGuard g(idx->device);
  unsigned long long h[4] = {0, 0, 0, 0};
  if (idx->stats.p && idx->searched) {
    // wait for the most recent search only (its last kernel recorded ev_done), then read the four counters
    CU_TRY(cudaEventSynchronize(idx->ev_done));
    CU_TRY(cudaMemcpyAsync(h, idx->stats.p, sizeof h, cudaMemcpyDeviceToHost, idx->side));
    CU_TRY(cudaStreamSynchronize(idx->side));
  }
  // with the hand-over on (default) every unproven query ends in the exhaustive scan at the latest, so none is left
  // unverified; with b2r_index_set_verify(idx, 0) the first pass's unproven queries are only counted
  idx->last.fallback_queries = idx->verify ? int64_t(h[0]) : 0;
  idx->last.candidates = int64_t(h[1]);
  idx->last.exhaustive_queries = int64_t(h[2]);
  idx->last.unverified = int64_t(h[3]);
  idx->last.fallback = idx->last.fallback_queries != 0 ? 1 : 0;
  *out = idx->last;
  return B2R_OK;
}

extern "C" int b2r_index_last_score_ms(b2r_index* idx, float* ms) {
  if (!idx || !ms) return fail(B2R_ERR_ARG, "null argument");
  std::lock_guard<std::mutex> lk(idx->mu);
  DeviceGuard g(idx->device);
  *ms = 0.f;
  if (!idx->have_times) return B2R_OK;
  const int slot = int((idx->nrec - 1) % b2r_index::kEvPool);
  CU_TRY(cudaEventSynchronize(idx->ev1[slot]));
  CU_TRY(cudaEventElapsedTime(ms, idx->ev0[slot], idx->ev1[slot]));
  return B2R_OK;
}

extern "C" int b2r_index_score_time(b2r_index* idx, int reset, double* total_ms, int64_t* nsearch) {
  if (!idx) return fail(B2R_ERR_ARG, "index is null");
  std::lock_guard<std::mutex> lk(idx->mu);
  DeviceGuard g(idx->device);
  const int64_t cnt = std::min<int64_t>(idx->nrec, b2r_index::kEvPool);
  double sum = 0.0;
  for (int64_t i = 0; i < cnt; ++i) {
    const int slot = int((idx->nrec - 1 - i) % b2r_index::kEvPool);
    float ms = 0.f;
    CU_TRY(cudaEventSynchronize(idx->ev1[slot]));
    CU_TRY(cudaEventElapsedTime(&ms, idx->ev0[slot], idx->ev1[slot]));
    sum += ms;
  }
  if (total_ms) *total_ms = sum;
  if (nsearch) *nsearch = cnt;
  if (reset) { idx->nrec = 0; idx->have_times = false; }
  return B2R_OK;
}

// --------------------------------------------------------------------------------------
// row-sharded search: one NCCL all-gather of the packed shard-local top-k, then the merge kernel
// --------------------------------------------------------------------------------------
// NCCL is bound at run time (dlopen): the library that is already loaded in the process wins (so a communicator
// created by the host framework's NCCL and one created here come from the same instance), else libnccl.so.2 from
// the loader path.  Only the five entry points below are used; their types are restated to keep nccl.h out of the
// build.
typedef struct { char internal[128]; } b2r_nccl_uid;
struct NcclApi {
  void* h = nullptr;
  int (*GetUniqueId)(b2r_nccl_uid*) = nullptr;
  int (*CommInitRank)(void**, int, b2r_nccl_uid, int) = nullptr;
  int (*CommDestroy)(void*) = nullptr;
  int (*AllGather)(const void*, void*, size_t, int, void*, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
};
static NcclApi* nccl_api() {
  static std::mutex mu;
  static NcclApi api;
  static bool tried = false;
  std::lock_guard<std::mutex> lk(mu);
  if (!tried) {
    tried = true;
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (h) {
      api.GetUniqueId = reinterpret_cast<int (*)(b2r_nccl_uid*)>(dlsym(h, "ncclGetUniqueId"));
      api.CommInitRank = reinterpret_cast<int (*)(void**, int, b2r_nccl_uid, int)>(dlsym(h, "ncclCommInitRank"));
      api.CommDestroy = reinterpret_cast<int (*)(void*)>(dlsym(h, "ncclCommDestroy"));
      api.AllGather = reinterpret_cast<int (*)(const void*, void*, size_t, int, void*, cudaStream_t)>(dlsym(h, "ncclAllGather"));
      api.GetErrorString = reinterpret_cast<const char* (*)(int)>(dlsym(h, "ncclGetErrorString"));
      if (api.GetUniqueId && api.CommInitRank && api.CommDestroy && api.AllGather && api.GetErrorString) api.h = h;
    }
  }
  return api.h ? &api : nullptr;
}
constexpr int kNcclUint64 = 5;   // ncclDataType_t ncclUint64
constexpr int kNcclFloat32 = 7;  // ncclDataType_t ncclFloat32

static void comm_release(b2r_index* idx) {
  if (idx->comm && idx->comm_owned) {
    NcclApi* n = nccl_api();
    if (n) n->CommDestroy(idx->comm);
  }
  idx->comm = nullptr;
  idx->comm_owned = false;
  idx->rank = 0;
  idx->world = 1;
  idx->row_offset = 0;
}

extern "C" int b2r_nccl_unique_id(void* out128) {
  if (!out128) return fail(B2R_ERR_ARG, "out is null");
  NcclApi* n = nccl_api();
  if (!n) return fail(B2R_ERR_STATE, "the NCCL library (libnccl.so.2) could not be loaded");
  b2r_nccl_uid id;
  const int r = n->GetUniqueId(&id);
  if (r != 0) return fail(B2R_ERR_CUDA, "ncclGetUniqueId failed: %s", n->GetErrorString(r));
  memcpy(out128, &id, sizeof id);
  return B2R_OK;
}

static int comm_check(b2r_index* idx, int rank, int world, int64_t row_offset) {
  if (!idx) return fail(B2R_ERR_ARG, "index is null");
  if (world < 1 || world > kFinMaxSlots || rank < 0 || rank >= world) return fail(B2R_ERR_ARG, "bad rank %d of %d", rank, world);
  if (row_offset < 0) return fail(B2R_ERR_ARG, "negative row offset");
  return B2R_OK;
}

extern "C" int b2r_index_set_comm(b2r_index* idx, void* nccl_comm, int rank, int world, int64_t row_offset) {
  RC_TRY(comm_check(idx, rank, world, row_offset));
  if (world > 1 && !nccl_comm) return fail(B2R_ERR_ARG, "nccl_comm is null");
  if (world > 1 && !nccl_api()) return fail(B2R_ERR_STATE, "the NCCL library (libnccl.so.2) could not be loaded");
  std::lock_guard<std::mutex> lk(idx->mu);
  comm_release(idx);
  idx->comm = world > 1 ? nccl_comm : nullptr;
  idx->comm_owned = false;
  idx->rank = rank; idx->world = world; idx->row_offset = row_offset;
  return B2R_OK;
}

extern "C" int b2r_index_comm_init(b2r_index* idx, const void* unique_id128, int rank, int world, int64_t row_offset) {
  RC_TRY(comm_check(idx, rank, world, row_offset));
  if (world > 1 && !unique_id128) return fail(B2R_ERR_ARG, "unique id is null");
  std::lock_guard<std::mutex> lk(idx->mu);
  comm_release(idx);
  if (world > 1) {
    NcclApi* n = nccl_api();
    if (!n) return fail(B2R_ERR_STATE, "the NCCL library (libnccl.so.2) could not be loaded");
    DeviceGuard g(idx->device);
    b2r_nccl_uid id;
    memcpy(&id, unique_id128, sizeof id);
    void* comm = nullptr;
    const int r = n->CommInitRank(&comm, world, id, rank);
    if (r != 0) return fail(B2R_ERR_CUDA, "ncclCommInitRank failed: %s", n->GetErrorString(r));
    idx->comm = comm;
    idx->comm_owned = true;
  }
  idx->rank = rank; idx->world = world; idx->row_offset = row_offset;
  return B2R_OK;
}

extern "C" int b2r_index_search_sharded(b2r_index* idx, int64_t nq, const float* q, int64_t k, int normalize_queries,
                                        float* D, int64_t* I, void* stream) {
  RC_TRY(check_search_args(idx, nq, q, k));
  if (nq == 0) return B2R_OK;
  if (!D || !I) return fail(B2R_ERR_ARG, "D or I is null");
  std::lock_guard<std::mutex> lk(idx->mu);
  if (idx->world > 1 && !idx->comm) return fail(B2R_ERR_STATE, "no communicator: call b2r_index_set_comm or b2r_index_comm_init first");
  if (idx->ntotal + idx->row_offset > 0xfffffffeLL) return fail(B2R_ERR_LIMIT, "global ids must stay below 2^32-1");
  NcclApi* n = idx->world > 1 ? nccl_api() : nullptr;
  DeviceGuard g(idx->device);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const bool out_dev = is_device_ptr(D), q_dev = is_device_ptr(q);
  if (out_dev != is_device_ptr(I)) return fail(B2R_ERR_ARG, "D and I must both be host or both be device pointers");
  const int64_t mb = std::min(nq, kSearchBatch);
  // handle-owned, persistent: this rank's packed top-k and the gathered [world][batch*k] lists
  RC_TRY(idx->shard_local.ensure(size_t(mb) * k * 8));
  RC_TRY(idx->shard_all.ensure(size_t(idx->world) * mb * k * 8));
  if (!out_dev) {
    RC_TRY(idx->stage_d.ensure(size_t(mb) * k * 4));
    RC_TRY(idx->stage_i.ensure(size_t(mb) * k * 8));
  }
  uint64_t* local = static_cast<uint64_t*>(idx->shard_local.p);
  uint64_t* all = static_cast<uint64_t*>(idx->shard_all.p);
  for (int64_t q0 = 0; q0 < nq; q0 += kSearchBatch) {
    const int64_t m = std::min(kSearchBatch, nq - q0);
    const float* qb = q + q0 * idx->d;
    if (!q_dev && idx->world > 1 && m >= 16 * idx->world) {
      // Host queries: every rank holds the same rows, so each uploads only ITS 1/G of them over PCIe and the ranks
      // exchange the slices over NVLink (an in-place all-gather), instead of G full uploads from one host at once.
      const int64_t per = (m + idx->world - 1) / idx->world;
      RC_TRY(idx->stage_q.ensure(size_t(per) * idx->world * idx->d * 4));
      float* sq = static_cast<float*>(idx->stage_q.p);
      const int64_t r0 = std::min<int64_t>(m, per * idx->rank), r1 = std::min<int64_t>(m, r0 + per);
      if (r1 > r0)
        CU_TRY(cudaMemcpyAsync(sq + r0 * idx->d, qb + r0 * idx->d, size_t(r1 - r0) * idx->d * 4, cudaMemcpyHostToDevice, st));
      const int r = n->AllGather(sq + per * idx->rank * idx->d, sq, size_t(per) * idx->d, kNcclFloat32, idx->comm, st);
      if (r != 0) return fail(B2R_ERR_CUDA, "ncclAllGather (queries) failed: %s", n->GetErrorString(r));
      qb = sq;
    }
    RC_TRY(search_batch(idx, m, qb, k, normalize_queries, idx->row_offset, nullptr, nullptr, local, st, q0 == 0));
    const uint64_t* lists = local;
    if (idx->world > 1) {
      const int r = n->AllGather(local, all, size_t(m) * k, kNcclUint64, idx->comm, st);   // the single collective
      if (r != 0) return fail(B2R_ERR_CUDA, "ncclAllGather failed: %s", n->GetErrorString(r));
      lists = all;
    }
    float* Dd = out_dev ? D + q0 * k : static_cast<float*>(idx->stage_d.p);
    int64_t* Id = out_dev ? I + q0 * k : static_cast<int64_t*>(idx->stage_i.p);
    RC_TRY(merge_launch(idx->metric, idx->world, m, k, lists, m * k, Dd, Id, idx->device, st));
    ++idx->last.launches;
    if (!out_dev) {
      CU_TRY(cudaMemcpyAsync(D + q0 * k, Dd, size_t(m) * k * 4, cudaMemcpyDeviceToHost, st));
      CU_TRY(cudaMemcpyAsync(I + q0 * k, Id, size_t(m) * k * 8, cudaMemcpyDeviceToHost, st));
      CU_TRY(cudaStreamSynchronize(st));
    } else if (!q_dev) {
      CU_TRY(cudaStreamSynchronize(st));          // the caller's host query buffer is free again on return
    }
  }
  idx->searched = true;
  CU_TRY(cudaEventRecord(idx->ev_done, st));
  return B2R_OK;
}

// --------------------------------------------------------------------------------------
// stand-alone kernels
// --------------------------------------------------------------------------------------
extern "C" int b2r_normalize_l2(int64_t n, int64_t d, float* x, int device, void* stream) {
  if (n < 0 || d <= 0 || d > 65536) return fail(B2R_ERR_ARG, "bad shape n=%lld d=%lld", (long long)n, (long long)d);
  if (n == 0) return B2R_OK;
  if (!x) return fail(B2R_ERR_ARG, "x is null");
  if (b2r_device_count() <= 0) return fail(B2R_ERR_CUDA, "no CUDA device visible: this library has no CPU path");
  DeviceGuard g(device);
  if (!g.ok) return fail(B2R_ERR_CUDA, "cudaSetDevice(%d) failed", device);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const bool dev = is_device_ptr(x);
  float* xd = x;
  if (!dev) {
    CU_TRY(cudaMalloc(&xd, size_t(n) * d * 4));
    cudaError_t e = cudaMemcpyAsync(xd, x, size_t(n) * d * 4, cudaMemcpyHostToDevice, st);
    if (e != cudaSuccess) { cudaFree(xd); return fail(B2R_ERR_CUDA, "H2D copy failed: %s", cudaGetErrorString(e)); }
  }
  int sm = 148;
  cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, device);
  const int blocks = int(std::min<int64_t>((n + 7) / 8, int64_t(sm) * 16));
  normalize_inplace_kernel<<<blocks, 256, 0, st>>>(xd, n, int(d));
  cudaError_t e = cudaGetLastError();
  if (e == cudaSuccess && !dev) {
    e = cudaMemcpyAsync(x, xd, size_t(n) * d * 4, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  }
  if (!dev) cudaFree(xd);
  if (e != cudaSuccess) return fail(B2R_ERR_CUDA, "normalize_l2 failed: %s", cudaGetErrorString(e));
  return B2R_OK;
}

extern "C" int b2r_cosine_rows(int64_t n, int64_t m, int64_t d, const float* resp, const float* docs, float* cosout,
                               float* maxcos, int device, void* stream) {
  if (n < 0 || m < 0 || d <= 0) return fail(B2R_ERR_ARG, "bad shape n=%lld m=%lld d=%lld", (long long)n, (long long)m, (long long)d);
  if (n == 0) return B2R_OK;
  if (!resp || !maxcos || (m > 0 && !docs)) return fail(B2R_ERR_ARG, "null argument");
  if (b2r_device_count() <= 0) return fail(B2R_ERR_CUDA, "no CUDA device visible: this library has no CPU path");
  DeviceGuard g(device);
  if (!g.ok) return fail(B2R_ERR_CUDA, "cudaSetDevice(%d) failed", device);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const bool dev = is_device_ptr(resp);
  if (m > 0 && dev != is_device_ptr(docs)) return fail(B2R_ERR_ARG, "resp and docs must both be host or both be device pointers");
  if (dev != is_device_ptr(maxcos) || (cosout && dev != is_device_ptr(cosout)))
    return fail(B2R_ERR_ARG, "outputs must live where the inputs live");
  const size_t br = size_t(n) * d * 4, bd = size_t(n) * m * d * 4, bc = size_t(n) * m * 4, bm = size_t(n) * 4;
  char* ws = nullptr;
  // layout: [maxord n*4][resp][docs][cos][maxcos] (the last four only for host callers)
  const size_t total = bm + (dev ? 0 : br + bd + bc + bm);
  CU_TRY(cudaMalloc(&ws, std::max<size_t>(total, 16)));
  uint32_t* maxord = reinterpret_cast<uint32_t*>(ws);
  const float* r_d = resp; const float* d_d = docs; float* c_d = cosout; float* m_d = maxcos;
  cudaError_t e = cudaMemsetAsync(maxord, 0, bm, st);
  if (!dev) {
    float* rr = reinterpret_cast<float*>(ws + bm);
    float* dd = reinterpret_cast<float*>(ws + bm + br);
    c_d = reinterpret_cast<float*>(ws + bm + br + bd);
    m_d = reinterpret_cast<float*>(ws + bm + br + bd + bc);
    if (e == cudaSuccess) e = cudaMemcpyAsync(rr, resp, br, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess && m > 0) e = cudaMemcpyAsync(dd, docs, bd, cudaMemcpyHostToDevice, st);
    r_d = rr; d_d = dd;
  }
  if (e == cudaSuccess && m > 0) {
    int sm = 148;
    cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, device);
    const int blocks = int(std::min<int64_t>((n * m + 7) / 8, int64_t(sm) * 16));
    cosine_rows_kernel<<<blocks, 256, 0, st>>>(r_d, d_d, n, m, int(d), (dev && !cosout) ? nullptr : c_d, maxord);
    e = cudaGetLastError();
  }
  if (e == cudaSuccess) {
    finish_maxcos_kernel<<<int((n + 255) / 256), 256, 0, st>>>(maxord, n, m_d);
    e = cudaGetLastError();
  }
  if (e == cudaSuccess && !dev) {
    if (cosout && m > 0) e = cudaMemcpyAsync(cosout, c_d, bc, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(maxcos, m_d, bm, cudaMemcpyDeviceToHost, st);
  }
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  cudaFree(ws);
  if (e != cudaSuccess) return fail(B2R_ERR_CUDA, "cosine_rows failed: %s", cudaGetErrorString(e));
  return B2R_OK;
}

// --------------------------------------------------------------------------------------
// allocation-free reward cosine
// --------------------------------------------------------------------------------------
// Who needs this: the reward scoring sits inside the PPO batch loop (reference :297-302), next to a policy
// whose framework has usually taken all free device memory into its own caching pool.  In such a process a
// raw cudaMalloc can fail although the pool has room, and cudaMalloc / cudaFree also serialise the whole
// device, stalling the policy's own streams.  Serving loops that forbid allocation after warm-up have the
// same constraint.  b2r_cosine_rows allocates a small workspace per call; this variant does not: its only
// scratch (one ordered-float running maximum per sample) is the fixed array below, resident on every GPU
// that loads the library until the process exits (256 KiB).
static __device__ __align__(256) uint32_t g_noalloc_maxord[B2R_NOALLOC_MAX_SAMPLES];

// g_noalloc_maxord is a host-side stand-in in host code, not a device address: the device address is looked
// up once per device (with that device current) and cached here.
static std::mutex g_noalloc_mu;
static uint32_t* g_noalloc_dev_ptr[64] = {};

extern "C" int b2r_cosine_rows_noalloc(int64_t n, int64_t m, int64_t d, const float* resp, const float* docs,
                                       float* cosout, float* maxcos, int device, void* stream) {
  if (n < 0 || m < 0 || d <= 0) return fail(B2R_ERR_ARG, "bad shape n=%lld m=%lld d=%lld", (long long)n, (long long)m, (long long)d);
  if (n > B2R_NOALLOC_MAX_SAMPLES)
    return fail(B2R_ERR_LIMIT, "n=%lld exceeds B2R_NOALLOC_MAX_SAMPLES=%d: the fixed scratch holds one entry per sample; "
                "split the batch or use b2r_cosine_rows", (long long)n, B2R_NOALLOC_MAX_SAMPLES);
  if (n == 0) return B2R_OK;
  if (!resp || !maxcos || (m > 0 && !docs)) return fail(B2R_ERR_ARG, "null argument");
  const int ndev = b2r_device_count();
  if (ndev <= 0) return fail(B2R_ERR_CUDA, "no CUDA device visible: this library has no CPU path");
  if (device < 0 || device >= ndev || device >= 64) return fail(B2R_ERR_ARG, "device %d out of range", device);
  {
    // every buffer must be device memory of `device` itself: a pointer owned by another GPU would fault in the
    // kernel (no peer mapping is set up here), so it is refused up front
    const void* ptrs[4] = {resp, maxcos, m > 0 ? docs : nullptr, cosout};
    for (const void* ptr : ptrs) {
      if (!ptr) continue;
      int owner = -1;
      if (!is_device_ptr(ptr, &owner))
        return fail(B2R_ERR_ARG, "b2r_cosine_rows_noalloc takes device pointers only (staging host buffers would need an allocation)");
      if (owner != device)
        return fail(B2R_ERR_ARG, "a buffer lives on device %d but the call names device %d", owner, device);
    }
  }
  DeviceGuard g(device);
  if (!g.ok) return fail(B2R_ERR_CUDA, "cudaSetDevice(%d) failed", device);
  uint32_t* scratch = nullptr;
  {
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_noalloc_dev_ptr[device]) {   // first call on this device: resolve the symbol's device address
      void* p = nullptr;
      cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_maxord);
      if (e != cudaSuccess || !p) {
        cudaGetLastError();
        return fail(B2R_ERR_CUDA, "cudaGetSymbolAddress(noalloc scratch) failed on device %d: %s", device, cudaGetErrorString(e));
      }
      g_noalloc_dev_ptr[device] = static_cast<uint32_t*>(p);
    }
    scratch = g_noalloc_dev_ptr[device];
  }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  // the array keeps the previous call's maxima: the entries this call uses must start at "no value" (0)
  CU_TRY(cudaMemsetAsync(scratch, 0, size_t(n) * sizeof(uint32_t), st));
  if (m > 0) {
    int sm = 148;
    cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, device);
    const int blocks = int(std::min<int64_t>((n * m + 7) / 8, int64_t(sm) * 16));
    cosine_rows_kernel<<<blocks, 256, 0, st>>>(resp, docs, n, m, int(d), cosout, scratch);
    CU_TRY(cudaGetLastError());
  }
  finish_maxcos_kernel<<<int((n + 255) / 256), 256, 0, st>>>(scratch, n, maxcos);
  CU_TRY(cudaGetLastError());
  return B2R_OK;
}

// --------------------------------------------------------------------------------------
// shard export / import (stored representation, no re-normalisation, no re-rounding)
// --------------------------------------------------------------------------------------
extern "C" int b2r_index_export_rows(b2r_index* idx, int64_t row0, int64_t n, void* out, void* stream) {
  if (!idx) return fail(B2R_ERR_ARG, "index is null");
  if (row0 < 0 || n < 0 || row0 + n > idx->ntotal) return fail(B2R_ERR_ARG, "rows [%lld, %lld) out of range (ntotal %lld)", (long long)row0, (long long)(row0 + n), (long long)idx->ntotal);
  if (n == 0) return B2R_OK;
  if (!out) return fail(B2R_ERR_ARG, "out is null");
  std::lock_guard<std::mutex> lk(idx->mu);
  DeviceGuard g(idx->device);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const size_t esz = idx->elem_bytes(), w = size_t(idx->d) * esz;
  const char* src = static_cast<const char*>(idx->X) + size_t(row0) * idx->ld * esz;
  CU_TRY(cudaMemcpy2DAsync(out, w, src, size_t(idx->ld) * esz, w, size_t(n), cudaMemcpyDefault, st));
  if (!is_device_ptr(out)) CU_TRY(cudaStreamSynchronize(st));
  return B2R_OK;
}

extern "C" int b2r_index_import_rows(b2r_index* idx, int64_t n, const void* rows, void* stream) {
  if (!idx) return fail(B2R_ERR_ARG, "index is null");
  if (n < 0) return fail(B2R_ERR_ARG, "negative row count");
  if (n == 0) return B2R_OK;
  if (!rows) return fail(B2R_ERR_ARG, "rows is null");
  if (idx->ntotal + n > 0xfffffffeLL) return fail(B2R_ERR_LIMIT, "a shard holds at most 2^32-2 rows");
  std::lock_guard<std::mutex> lk(idx->mu);
  DeviceGuard g(idx->device);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  RC_TRY(grow(idx, idx->ntotal + n, st));
  const size_t esz = idx->elem_bytes(), w = size_t(idx->d) * esz, pitch = size_t(idx->ld) * esz;
  char* dst = static_cast<char*>(idx->X) + size_t(idx->ntotal) * pitch;
  if (pitch != w) CU_TRY(cudaMemsetAsync(dst, 0, size_t(n) * pitch, st));   // zero padding of every row
  CU_TRY(cudaMemcpy2DAsync(dst, pitch, rows, w, w, size_t(n), cudaMemcpyDefault, st));
  const int blocks = int(std::min<int64_t>((n + 7) / 8, int64_t(idx->sm_count) * 16));
  float* n2 = idx->norm2 ? idx->norm2 + idx->ntotal : nullptr;
  if (idx->storage == B2R_STORE_BF16)
    row_norms_kernel<__nv_bfloat16><<<blocks, 256, 0, st>>>(reinterpret_cast<const __nv_bfloat16*>(dst), n, int(idx->d), idx->ld, n2, idx->maxnorm2);
  else
    row_norms_kernel<float><<<blocks, 256, 0, st>>>(reinterpret_cast<const float*>(dst), n, int(idx->d), idx->ld, n2, idx->maxnorm2);
  if (idx->shadow)
    shadow_rows_kernel<<<blocks, 256, 0, st>>>(reinterpret_cast<const float*>(dst), n, idx->ld,
                                               static_cast<__nv_bfloat16*>(idx->Xs) + idx->ntotal * idx->ld);
  CU_TRY(cudaGetLastError());
  if (!is_device_ptr(rows)) CU_TRY(cudaStreamSynchronize(st));   // the caller may reuse its host buffer
  idx->ntotal += n;
  return B2R_OK;
}
