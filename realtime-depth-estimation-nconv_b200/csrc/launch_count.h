// Per-process launch accounting for libnconv_b200.so.
//  * count: how many kernels this library enqueued (bench.py reports it as gpu_launches);
//  * optional per-launch CUDA-event timing on the launching stream (bench.py's roofline leg):
//    when enabled, every launcher brackets its kernel with two events and records a label plus
//    the launch's algorithmic flops and bytes.  Off by default; no cost beyond one atomic add.
#pragma once
#include <cuda_runtime.h>
#include <atomic>

namespace nconv {
extern std::atomic<unsigned long long> g_launch_count;
extern std::atomic<int> g_profile_on;

inline void count_launch(unsigned long long n = 1) { g_launch_count.fetch_add(n, std::memory_order_relaxed); }

// implemented in elementwise.cu
void profile_open(cudaStream_t st, const char* label, double flops, double bytes, void** token);
void profile_close(cudaStream_t st, void* token);

// RAII bracket around exactly one kernel launch.
struct LaunchScope {
  cudaStream_t st;
  void* token;
  LaunchScope(cudaStream_t s, const char* label, double flops, double bytes) : st(s), token(nullptr) {
    count_launch();
    if (g_profile_on.load(std::memory_order_relaxed)) profile_open(st, label, flops, bytes, &token);
  }
  ~LaunchScope() {
    if (token) profile_close(st, token);
  }
};
}  // namespace nconv
