// score_umma.cuh — kernel (b), tensor-core variant (placeholder until the tcgen05 kernel lands).
#pragma once
#include <string>
#include "common.cuh"
#include "finalize.cuh"

namespace b2r {

struct UmmaState {
  void release() {}
  void invalidate() {}
};

struct UmmaCall {
  const void* X; int64_t n; int ld; int d; int storage;
  const void* Qh; const float* Qf; int nq; int kp;
  const float* norm2; uint32_t* thr; int sm_count;
};

inline bool umma_supported(int, int, int, int64_t) { return false; }
inline int umma_launch(UmmaState&, const UmmaCall&, cudaStream_t, FinalizeParams&, int*, std::string& err) {
  err = "tcgen05 engine not built";
  return -5;
}

}  // namespace b2r
