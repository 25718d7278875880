// Optional per-kernel timing with CUDA events on the launching stream (off by default).
// bench.py switches it on around its timed region to get each kernel's average launch duration
// and the number of launches; see te_profile_enable / te_profile_read in include/te_b200.h.
#pragma once
#include <cuda_runtime.h>

namespace te {

struct ProfScope {
  ProfScope(const char* name, cudaStream_t st);
  ~ProfScope();
  int slot_;
  cudaStream_t st_;
};

}  // namespace te

#define TE_PROF(name, st) ::te::ProfScope _te_prof_scope_(name, st)
