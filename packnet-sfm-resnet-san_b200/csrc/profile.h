// profile.h -- launch accounting and optional per-kernel CUDA-event timing (off by default).
#pragma once
#include <cuda_runtime.h>

namespace psfm {
enum ProfTag { kTagMean = 0, kTagFwd = 1, kTagBwd = 2, kTagFinish = 3, kTagFused = 4, kTagOther = 5 };
void count_launches(int n);
// records an event pair around the enclosed launches when profiling is enabled
struct ProfScope {
  ProfScope(cudaStream_t st, int tag, int launches);
  ~ProfScope();
  cudaStream_t st_;
  int slot_;
};
}  // namespace psfm
