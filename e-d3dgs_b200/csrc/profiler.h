// Optional per-kernel timing with CUDA events on the launching stream (off by default).
// bench.py switches it on for its roofline pass: ed3dgs_profile_enable(1), run K steps,
// ed3dgs_profile_report() -> "name total_ms launches" lines. When off, a ProfScope is two
// predictable branches.
#pragma once
#include <cuda_runtime.h>

namespace ed3dgs {

bool prof_enabled();
void prof_push(const char* name, cudaEvent_t start, cudaEvent_t stop);
// Device counters {pairs walked by render_fwd, pairs walked by blend_bwd} while profiling is on, else nullptr.
unsigned long long* prof_pair_counters();

struct ProfScope {
  cudaEvent_t start_ = nullptr, stop_ = nullptr;
  cudaStream_t stream_;
  const char* name_;
  ProfScope(const char* name, cudaStream_t stream) : stream_(stream), name_(name) {
    if (prof_enabled()) {
      cudaEventCreate(&start_);
      cudaEventCreate(&stop_);
      cudaEventRecord(start_, stream_);
    }
  }
  ~ProfScope() {
    if (start_) {
      cudaEventRecord(stop_, stream_);
      prof_push(name_, start_, stop_);
    }
  }
};

}  // namespace ed3dgs
