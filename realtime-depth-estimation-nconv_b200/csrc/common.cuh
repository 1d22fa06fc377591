// Shared helpers for the sm_100a NConv kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define NCONV_MAX_DEVICES 16

// Launchers never throw and never allocate: they return the cudaError_t of the enqueue.
#define NCONV_RETURN_IF_ERR(expr)            \
  do {                                       \
    cudaError_t _e = (expr);                 \
    if (_e != cudaSuccess) return (int)_e;   \
  } while (0)

// Error codes above the CUDA range for argument problems detected by a launcher.
enum { NCONV_ERR_BAD_ARG = 100001, NCONV_ERR_UNSUPPORTED = 100002, NCONV_ERR_WORKSPACE = 100003, NCONV_ERR_STATIC_SCRATCH = 100004 };

namespace nconv {

__host__ __device__ __forceinline__ int ceil_div(int a, int b) { return (a + b - 1) / b; }

// Nearest-neighbour source index, restating ATen's `nearest_idx`
// (aten/src/ATen/native/cpu/UpSampleKernel.cpp; used by F.interpolate(mode='nearest') at
// /root/reference/models/step1.py:78-89): identity when sizes match, >>1 for exact 2x,
// otherwise min(floor(dst * (float)in/out), in-1).
struct NearestMap {
  int mode;     // 0: same, 1: exact x2, 2: general
  float scale;  // (float)in / out
  int in_size;
  __host__ static NearestMap make(int in_size, int out_size) {
    NearestMap m;
    m.in_size = in_size;
    m.scale = (float)in_size / (float)out_size;
    m.mode = (in_size == out_size) ? 0 : (out_size == 2 * in_size ? 1 : 2);
    return m;
  }
  __device__ __forceinline__ int operator()(int dst) const {
    if (mode == 0) return dst;
    if (mode == 1) return dst >> 1;
    int s = (int)floorf((float)dst * scale);
    return s < in_size - 1 ? s : in_size - 1;
  }
};

// max with ATen max_pool2d semantics: `(val > maxval) || isnan(val)` (first max wins on
// ties, NaN propagates) — /root/reference/models/step1.py:62-74 via F.max_pool2d.
__device__ __forceinline__ float pool_max(float m, float v) { return (v > m || v != v) ? v : m; }

// Same value as a chain of pool_max updates over a window (max, NaN if any element is NaN), one FMNMX.NAN each.
__device__ __forceinline__ float max_nan(float a, float b) {
  float d;
  asm("max.NaN.f32 %0, %1, %2;" : "=f"(d) : "f"(a), "f"(b));
  return d;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
  return v;
}

// softplus with beta=10 and ATen's threshold=20: x if 10x > 20 else log1p(exp(10x))/10
// (/root/reference/models/step1.py:207).
__device__ __forceinline__ float softplus_beta10(float x) {
  float bx = x * 10.0f;
  return bx > 20.0f ? x : log1pf(expf(bx)) / 10.0f;
}

// Packed fp32 FMA (sm_100 FFMA2): acc.{lo,hi} += w.{lo,hi} * s.  ptxas folds the {s, s} pack into
// the instruction's scalar-broadcast operand form (`FFMA2 Rd, Rw.F32x2.HI_LO, Rs.F32, Rd`), so no
// extra register or MOV is spent on it.
// Ampere-style asynchronous global -> shared copies; src-size 0 zero-fills the destination (out-of-image elements).
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc, bool ok) {
  const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  const int n = ok ? 16 : 0;
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(d), "l"(gsrc), "r"(n) : "memory");
}
__device__ __forceinline__ void cp_async8(void* smem_dst, const void* gsrc, bool ok) {
  const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  const int n = ok ? 8 : 0;
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;" ::"r"(d), "l"(gsrc), "r"(n) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

__device__ __forceinline__ void cp_async4(void* smem_dst, const void* gsrc, bool ok) {
  const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  const int n = ok ? 4 : 0;
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(d), "l"(gsrc), "r"(n) : "memory");
}

typedef unsigned long long u64;
__device__ __forceinline__ void ffma2(u64& acc, u64 w, float s) {
  asm("{\n\t.reg .b64 t;\n\tmov.b64 t, {%2, %2};\n\tfma.rn.f32x2 %0, %1, t, %0;\n\t}" : "+l"(acc) : "l"(w), "f"(s));
}
__device__ __forceinline__ float lo32f(u64 v) { return __uint_as_float((unsigned)(v & 0xffffffffull)); }
__device__ __forceinline__ float hi32f(u64 v) { return __uint_as_float((unsigned)(v >> 32)); }

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

}  // namespace nconv
