// Training-mode kernels of the RGB-guided stack (SETP2_BP_TRAIN, /root/reference/models/step2.py:22-77
// run under train_step2.py:60-66, i.e. model.train(): BatchNorm2d normalises with BATCH statistics and
// autograd differentiates every Conv2d / ConvTranspose2d / BatchNorm2d / ReLU of the stack).  fp32 NCHW.
//
//   bn_train_stats / bn_train_finalize / affine_act   BatchNorm2d training forward (+ReLU, + residual)
//   bn_bwd_reduce / bn_bwd_apply                      BatchNorm2d (+ReLU) backward
//   relu_bias_bwd                                     ConvBlock's ReLU(conv + bias) backward: masks the
//                                                     gradient and reduces the bias gradient
//   channel_sum                                       bias gradient of a plain convolution
//   conv_wgrad<K,S>                                   weight gradient (also ConvTranspose2d's, roles swapped)
//   conv_gather<MODE,K,S>                             data gradient of the stride-2 convolutions
//                                                     (transposed gather) and of ConvTranspose2d(4,2,1)
//                                                     (a strided 4x4 convolution of the gradient)
//   wflip                                             W[co][ci][u][v] -> W'[ci][co][K-1-u][K-1-v]: the data
//                                                     gradient of a stride-1 "same" convolution is the
//                                                     forward kernel (conv_simt.cu) run on W'
//
// Per-channel reductions accumulate in double (per-thread fp32 partials over <= 64 values, combined in
// fp64 through shared memory and one atomicAdd(double) per CTA), so BatchNorm statistics of the
// un-normalised RGB activations (mean >> std) do not lose digits and results do not depend on the grid.
#include <mutex>
#include "common.cuh"
#include <cstdint>
#include <type_traits>
#include "nconv_b200.h"
#include "launch_count.h"
#include <cstdio>
#include <cstdlib>

namespace nconv {

constexpr int RED_THREADS = 256;

__device__ __forceinline__ double block_sum_double(double v, double* sh) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
  __syncthreads();
  if (lane == 0) sh[warp] = v;
  __syncthreads();
  double t = 0.0;
  if (warp == 0) {
    t = (lane < RED_THREADS / 32) ? sh[lane] : 0.0;
#pragma unroll
    for (int d = 4; d > 0; d >>= 1) t += __shfl_xor_sync(0xffffffffu, t, d);
  }
  return t;   // valid in thread 0
}

// ---- BatchNorm training forward ---------------------------------------------------------------------
// acc[c][0] += sum(x - pivot[c]),  acc[c][1] += sum((x - pivot[c])^2); pivot = x[0, c, 0, 0] keeps the
// one-pass variance well conditioned.  grid (chunks, C, B).
__global__ void __launch_bounds__(RED_THREADS) bn_stats_kernel(const float* __restrict__ x, double* __restrict__ acc,
                                                               int C, int HW) {
  __shared__ double sh[RED_THREADS / 32];
  const int c = blockIdx.y, b = blockIdx.z;
  const float pivot = __ldg(x + (size_t)c * HW);
  const float* p = x + ((size_t)b * C + c) * HW;
  double s1 = 0.0, s2 = 0.0;
  // (scalar, strided order kept on purpose: the batch statistics feed the forward activations, and a training step of this
  //  stack is chaotic at the 1e-3 level -- a different summation order here re-rolls every ReLU mask decision near zero)
  const int per = (HW + gridDim.x - 1) / gridDim.x;
  const int lo = blockIdx.x * per, hi = min(HW, lo + per);
  for (int i0 = lo + threadIdx.x; i0 < hi; i0 += RED_THREADS * 32) {
    float f1 = 0.f, f2 = 0.f;
#pragma unroll 4
    for (int k = 0; k < 32; ++k) {
      const int i = i0 + k * RED_THREADS;
      if (i < hi) { const float d = __ldg(p + i) - pivot; f1 += d; f2 = fmaf(d, d, f2); }
    }
    s1 += f1; s2 += f2;
  }
  s1 = block_sum_double(s1, sh);
  s2 = block_sum_double(s2, sh);
  if (threadIdx.x == 0) { atomicAdd(acc + 2 * c, s1); atomicAdd(acc + 2 * c + 1, s2); }
}

// mean/var from the sums; scale/shift for the apply pass; save mean and rstd for backward; update the
// running statistics exactly as nn.BatchNorm2d does (momentum, unbiased variance).
__global__ void bn_finalize_kernel(const double* __restrict__ acc, const float* __restrict__ x, const float* __restrict__ gamma,
                                   const float* __restrict__ beta, float* __restrict__ running_mean,
                                   float* __restrict__ running_var, float* __restrict__ scale, float* __restrict__ shift,
                                   float* __restrict__ save_mean, float* __restrict__ save_rstd, int C, int HW, double n,
                                   float eps, float momentum) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const double pivot = (double)__ldg(x + (size_t)c * HW);
  const double m1 = acc[2 * c] / n, m2 = acc[2 * c + 1] / n;
  const double mean = pivot + m1;
  double var = m2 - m1 * m1;
  if (var < 0.0) var = 0.0;
  const float rstd = (float)(1.0 / sqrt(var + (double)eps));
  const float g = gamma ? gamma[c] : 1.f, bt = beta ? beta[c] : 0.f;
  scale[c] = g * rstd;
  shift[c] = bt - (float)mean * g * rstd;
  save_mean[c] = (float)mean;
  save_rstd[c] = rstd;
  if (running_mean) {
    const double unbiased = n > 1.0 ? var * n / (n - 1.0) : var;
    running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * (float)mean;
    running_var[c] = (1.f - momentum) * running_var[c] + momentum * (float)unbiased;
  }
}

// y = act(x * scale[c] + shift[c]) (+ residual); in place allowed (y == x).  grid (chunks, C, B): a CTA streams one slice of
// one (b, c) plane, so the channel constants are registers and there is no per-element index arithmetic; 128-bit
// accesses when the plane size allows (VEC = 4).
template <int VEC>
__global__ void __launch_bounds__(256) affine_act_kernel(const float* __restrict__ x, const float* __restrict__ scale,
                                                         const float* __restrict__ shift, const float* __restrict__ residual,
                                                         float* __restrict__ y, int C, int HW, int per, int relu) {
  const int c = blockIdx.y, b = blockIdx.z;
  const float sc = __ldg(scale + c), sf = __ldg(shift + c);
  const size_t base = ((size_t)b * C + c) * HW;
  const int lo = blockIdx.x * per, hi = min(HW, lo + per);
  for (int i = lo + VEC * threadIdx.x; i < hi; i += VEC * 256) {
    float v[VEC], r[VEC];
    if (VEC == 4) {
      *reinterpret_cast<float4*>(v) = *reinterpret_cast<const float4*>(x + base + i);
      if (residual) *reinterpret_cast<float4*>(r) = __ldg(reinterpret_cast<const float4*>(residual + base + i));
    } else {
      v[0] = x[base + i];
      if (residual) r[0] = __ldg(residual + base + i);
    }
#pragma unroll
    for (int k = 0; k < VEC; ++k) {
      v[k] = fmaf(v[k], sc, sf);
      if (relu) v[k] = fmaxf(v[k], 0.f);
      if (residual) v[k] += r[k];
    }
    if (VEC == 4) *reinterpret_cast<float4*>(y + base + i) = *reinterpret_cast<float4*>(v);
    else y[base + i] = v[0];
  }
}

// slices per plane for the streaming (non-reducing) kernels: >= 4 CTAs per SM in flight, a slice a multiple of 4 elements
static inline int stream_chunks(int HW, int C, int B, int* per) {
  int want = (148 * 8 + C * B - 1) / (C * B);
  const int maxc = (HW + 4095) / 4096;
  if (want > maxc) want = maxc;
  if (want < 1) want = 1;
  int p = (HW + want - 1) / want;
  p = (p + 3) / 4 * 4;
  *per = p;
  return (HW + p - 1) / p;
}

// ---- BatchNorm (+ReLU) backward ---------------------------------------------------------------------
// z: the BatchNorm input saved by forward.  g = gy * [z*scale+shift > 0] (ReLU mask, if relu).
// acc[c][0] += sum g, acc[c][1] += sum g * xhat.
__global__ void __launch_bounds__(RED_THREADS) bn_bwd_reduce_kernel(const float* __restrict__ gy, const float* __restrict__ z,
                                                                    const float* __restrict__ scale, const float* __restrict__ shift,
                                                                    const float* __restrict__ mean, const float* __restrict__ rstd,
                                                                    double* __restrict__ acc, int C, int HW, int relu) {
  __shared__ double sh[RED_THREADS / 32];
  const int c = blockIdx.y, b = blockIdx.z;
  const float sc = scale[c], sf = shift[c], mu = mean[c], rs = rstd[c];
  const size_t base = ((size_t)b * C + c) * HW;
  double s1 = 0.0, s2 = 0.0;
  const int per = ((HW + gridDim.x - 1) / gridDim.x + 3) & ~3;
  const int lo = blockIdx.x * per, hi = min(HW, lo + per);
  if ((HW & 3) == 0) {
    // 128-bit loads; fp32 partials over 8 iterations (32 values per thread), then into the fp64 sums
    for (int i0 = lo + 4 * threadIdx.x; i0 < hi; i0 += RED_THREADS * 32) {
      float f1 = 0.f, f2 = 0.f;
#pragma unroll 4
      for (int k = 0; k < 8; ++k) {
        const int i = i0 + k * RED_THREADS * 4;
        if (i < hi) {
          const float4 zv = __ldg(reinterpret_cast<const float4*>(z + base + i));
          float4 g = __ldg(reinterpret_cast<const float4*>(gy + base + i));
          if (relu) {
            if (!(fmaf(zv.x, sc, sf) > 0.f)) g.x = 0.f;
            if (!(fmaf(zv.y, sc, sf) > 0.f)) g.y = 0.f;
            if (!(fmaf(zv.z, sc, sf) > 0.f)) g.z = 0.f;
            if (!(fmaf(zv.w, sc, sf) > 0.f)) g.w = 0.f;
          }
          f1 += (g.x + g.y) + (g.z + g.w);
          f2 = fmaf(g.x, (zv.x - mu) * rs, f2); f2 = fmaf(g.y, (zv.y - mu) * rs, f2);
          f2 = fmaf(g.z, (zv.z - mu) * rs, f2); f2 = fmaf(g.w, (zv.w - mu) * rs, f2);
        }
      }
      s1 += f1; s2 += f2;
    }
  } else
  for (int i0 = lo + threadIdx.x; i0 < hi; i0 += RED_THREADS * 32) {
    float f1 = 0.f, f2 = 0.f;
#pragma unroll 4
    for (int k = 0; k < 32; ++k) {
      const int i = i0 + k * RED_THREADS;
      if (i < hi) {
        const float zv = __ldg(z + base + i);
        float g = __ldg(gy + base + i);
        if (relu && !(fmaf(zv, sc, sf) > 0.f)) g = 0.f;
        f1 += g; f2 = fmaf(g, (zv - mu) * rs, f2);
      }
    }
    s1 += f1; s2 += f2;
  }
  s1 = block_sum_double(s1, sh);
  s2 = block_sum_double(s2, sh);
  if (threadIdx.x == 0) { atomicAdd(acc + 2 * c, s1); atomicAdd(acc + 2 * c + 1, s2); }
}

// gz = gamma*rstd * (g - sum_g/n - xhat * sum_gxhat/n); also emits ggamma = sum_gxhat, gbeta = sum_g once.
// grid (chunks, C, B) as affine_act_kernel: the per-channel constants (two fp64 divisions among them) once per CTA.
template <int VEC>
__global__ void __launch_bounds__(256) bn_bwd_apply_kernel(const float* __restrict__ gy, const float* __restrict__ z,
                                                           const float* __restrict__ scale, const float* __restrict__ shift,
                                                           const float* __restrict__ mean, const float* __restrict__ rstd,
                                                           const float* __restrict__ gamma, const double* __restrict__ acc,
                                                           float* __restrict__ gz, float* __restrict__ ggamma,
                                                           float* __restrict__ gbeta, int C, int HW, int per, double n,
                                                           int relu) {
  const int c = blockIdx.y, b = blockIdx.z;
  if (blockIdx.x == 0 && b == 0 && threadIdx.x == 0) {
    if (gbeta) gbeta[c] = (float)acc[2 * c];
    if (ggamma) ggamma[c] = (float)acc[2 * c + 1];
  }
  const float sc = __ldg(scale + c), sf = __ldg(shift + c), mu = __ldg(mean + c), rs = __ldg(rstd + c);
  const float m1 = (float)(acc[2 * c] / n), m2 = (float)(acc[2 * c + 1] / n);
  const float k = (gamma ? __ldg(gamma + c) : 1.f) * rs;
  const size_t base = ((size_t)b * C + c) * HW;
  const int lo = blockIdx.x * per, hi = min(HW, lo + per);
  for (int i = lo + VEC * threadIdx.x; i < hi; i += VEC * 256) {
    float zv[VEC], g[VEC];
    if (VEC == 4) {
      *reinterpret_cast<float4*>(zv) = __ldg(reinterpret_cast<const float4*>(z + base + i));
      *reinterpret_cast<float4*>(g) = __ldg(reinterpret_cast<const float4*>(gy + base + i));
    } else {
      zv[0] = __ldg(z + base + i); g[0] = __ldg(gy + base + i);
    }
#pragma unroll
    for (int e = 0; e < VEC; ++e) {
      if (relu && !(fmaf(zv[e], sc, sf) > 0.f)) g[e] = 0.f;
      const float xh = (zv[e] - mu) * rs;
      g[e] = k * (g[e] - m1 - xh * m2);
    }
    if (VEC == 4) *reinterpret_cast<float4*>(gz + base + i) = *reinterpret_cast<float4*>(g);
    else gz[base + i] = g[0];
  }
}

// ---- ReLU(conv + bias) backward: g = gy * [y > 0]; gbias[c] = sum g --------------------------------
__global__ void __launch_bounds__(RED_THREADS) relu_bias_bwd_kernel(const float* __restrict__ gy, const float* __restrict__ y,
                                                                    float* __restrict__ g, double* __restrict__ acc, int C,
                                                                    int HW, int relu) {
  __shared__ double sh[RED_THREADS / 32];
  const int c = blockIdx.y, b = blockIdx.z;
  const size_t base = ((size_t)b * C + c) * HW;
  double s1 = 0.0;
  const int per = ((HW + gridDim.x - 1) / gridDim.x + 3) & ~3;
  const int lo = blockIdx.x * per, hi = min(HW, lo + per);
  if ((HW & 3) == 0) {
    for (int i0 = lo + 4 * threadIdx.x; i0 < hi; i0 += RED_THREADS * 32) {
      float f1 = 0.f;
#pragma unroll 4
      for (int k = 0; k < 8; ++k) {
        const int i = i0 + k * RED_THREADS * 4;
        if (i < hi) {
          float4 v = __ldg(reinterpret_cast<const float4*>(gy + base + i));
          if (relu) {
            const float4 yv = __ldg(reinterpret_cast<const float4*>(y + base + i));
            if (!(yv.x > 0.f)) v.x = 0.f;
            if (!(yv.y > 0.f)) v.y = 0.f;
            if (!(yv.z > 0.f)) v.z = 0.f;
            if (!(yv.w > 0.f)) v.w = 0.f;
          }
          if (g) *reinterpret_cast<float4*>(g + base + i) = v;
          f1 += (v.x + v.y) + (v.z + v.w);
        }
      }
      s1 += f1;
    }
  } else
  for (int i0 = lo + threadIdx.x; i0 < hi; i0 += RED_THREADS * 32) {
    float f1 = 0.f;
#pragma unroll 4
    for (int k = 0; k < 32; ++k) {
      const int i = i0 + k * RED_THREADS;
      if (i < hi) {
        float v = __ldg(gy + base + i);
        if (relu && !(__ldg(y + base + i) > 0.f)) v = 0.f;
        if (g) g[base + i] = v;
        f1 += v;
      }
    }
    s1 += f1;
  }
  s1 = block_sum_double(s1, sh);
  if (threadIdx.x == 0) atomicAdd(acc + c, s1);
}

__global__ void double_to_float_kernel(const double* __restrict__ a, float* __restrict__ o, int n, int stride) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) o[i] = (float)a[(size_t)i * stride];
}

static inline int red_chunks(int HW, int C, int B) {
  // enough CTAs to fill the machine, each with >= 8K elements
  int want = (148 * 8 + C * B - 1) / (C * B);
  int maxc = (HW + 8191) / 8192;
  int ch = want < maxc ? want : maxc;
  return ch < 1 ? 1 : ch;
}

}  // namespace nconv

using namespace nconv;

extern "C" size_t nconv_b200_train_ws_bytes(int C) { return (size_t)2 * C * sizeof(double); }

// BatchNorm2d training forward (+ReLU, + residual added after the activation, RGBEncoder :151-152).
// ws: nconv_b200_train_ws_bytes(C) of scratch.  scale/shift/save_mean/save_rstd: [C] outputs.
extern "C" int nconv_b200_bn_train_fwd(const float* x, const float* gamma, const float* beta, float* running_mean,
                                       float* running_var, const float* residual, float* y, float* scale, float* shift,
                                       float* save_mean, float* save_rstd, void* ws, int B, int C, int H, int W,
                                       float eps, float momentum, int relu, void* stream) {
  if (!x || !y || !scale || !shift || !save_mean || !save_rstd || !ws) return NCONV_ERR_BAD_ARG;
  if (B <= 0 || C <= 0 || H <= 0 || W <= 0) return NCONV_ERR_BAD_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  const int HW = H * W;
  const long long total = (long long)B * C * HW;
  NCONV_RETURN_IF_ERR(cudaMemsetAsync(ws, 0, nconv_b200_train_ws_bytes(C), st));
  {
    LaunchScope ls(st, "bn_train_stats", 3.0 * total, 4.0 * total);
    bn_stats_kernel<<<dim3(red_chunks(HW, C, B), C, B), RED_THREADS, 0, st>>>(x, (double*)ws, C, HW);
  }
  {
    LaunchScope ls(st, "bn_train_finalize", 0.0, 0.0);
    bn_finalize_kernel<<<(C + 127) / 128, 128, 0, st>>>((const double*)ws, x, gamma, beta, running_mean, running_var, scale,
                                                        shift, save_mean, save_rstd, C, HW, (double)B * HW, eps, momentum);
  }
  {
    LaunchScope ls(st, "bn_train_apply", 2.0 * total, 4.0 * total * (residual ? 3.0 : 2.0));
    int per = 0;
    const int chunks = stream_chunks(HW, C, B, &per);
    if (HW % 4 == 0) affine_act_kernel<4><<<dim3(chunks, C, B), 256, 0, st>>>(x, scale, shift, residual, y, C, HW, per, relu);
    else affine_act_kernel<1><<<dim3(chunks, C, B), 256, 0, st>>>(x, scale, shift, residual, y, C, HW, per, relu);
  }
  return (int)cudaGetLastError();
}

// BatchNorm2d (+ReLU) backward.  z = the BatchNorm input of forward; scale/shift/mean/rstd as saved.
extern "C" int nconv_b200_bn_train_bwd(const float* gy, const float* z, const float* scale, const float* shift,
                                       const float* mean, const float* rstd, const float* gamma, float* gz, float* ggamma,
                                       float* gbeta, void* ws, int B, int C, int H, int W, int relu, void* stream) {
  if (!gy || !z || !scale || !shift || !mean || !rstd || !gz || !ws) return NCONV_ERR_BAD_ARG;
  if (B <= 0 || C <= 0 || H <= 0 || W <= 0) return NCONV_ERR_BAD_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  const int HW = H * W;
  const long long total = (long long)B * C * HW;
  NCONV_RETURN_IF_ERR(cudaMemsetAsync(ws, 0, nconv_b200_train_ws_bytes(C), st));
  {
    LaunchScope ls(st, "bn_bwd_reduce", 5.0 * total, 8.0 * total);
    bn_bwd_reduce_kernel<<<dim3(red_chunks(HW, C, B), C, B), RED_THREADS, 0, st>>>(gy, z, scale, shift, mean, rstd,
                                                                                   (double*)ws, C, HW, relu);
  }
  {
    LaunchScope ls(st, "bn_bwd_apply", 8.0 * total, 12.0 * total);
    int per = 0;
    const int chunks = stream_chunks(HW, C, B, &per);
    if (HW % 4 == 0)
      bn_bwd_apply_kernel<4><<<dim3(chunks, C, B), 256, 0, st>>>(gy, z, scale, shift, mean, rstd, gamma, (const double*)ws, gz,
                                                                 ggamma, gbeta, C, HW, per, (double)B * HW, relu);
    else
      bn_bwd_apply_kernel<1><<<dim3(chunks, C, B), 256, 0, st>>>(gy, z, scale, shift, mean, rstd, gamma, (const double*)ws, gz,
                                                                 ggamma, gbeta, C, HW, per, (double)B * HW, relu);
  }
  return (int)cudaGetLastError();
}

// g = gy * [y > 0] (if relu; g may be null when only the channel sums are wanted), gbias[c] = sum g.
extern "C" int nconv_b200_relu_bias_bwd(const float* gy, const float* y, float* g, float* gbias, void* ws, int B, int C,
                                        int H, int W, int relu, void* stream) {
  if (!gy || !ws || (relu && !y)) return NCONV_ERR_BAD_ARG;
  if (B <= 0 || C <= 0 || H <= 0 || W <= 0) return NCONV_ERR_BAD_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  const int HW = H * W;
  const long long total = (long long)B * C * HW;
  NCONV_RETURN_IF_ERR(cudaMemsetAsync(ws, 0, nconv_b200_train_ws_bytes(C), st));
  {
    LaunchScope ls(st, "relu_bias_bwd", 2.0 * total, 4.0 * total * (g ? 3.0 : (relu ? 2.0 : 1.0)));
    relu_bias_bwd_kernel<<<dim3(red_chunks(HW, C, B), C, B), RED_THREADS, 0, st>>>(gy, y, g, (double*)ws, C, HW, relu);
  }
  if (gbias) {
    LaunchScope ls(st, "channel_sum_out", 0.0, 0.0);
    double_to_float_kernel<<<(C + 127) / 128, 128, 0, st>>>((const double*)ws, gbias, C, 1);
  }
  return (int)cudaGetLastError();
}

// ---- the same operation for callers that must not allocate ----------------------------------------------------------
// (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc.)  The per-channel fp64 sums live in a file-scope
// device array instead of a caller workspace: 2 doubles per channel, NCONV_B200_NOALLOC_MAX_C channels = 16 KB, held by
// every GPU that loads the library until the process exits.  In HOST code `g_noalloc_sums` is only a stand-in for the
// symbol, not a device address: its address is taken once per device with cudaGetSymbolAddress (that device current)
// and that pointer is what the kernels get.  All calls share the array and it keeps the previous call's sums: it is
// cleared (memset on the caller's stream, no allocation) at the start of every call; two calls that overlap on the GPU
// (different streams, no dependency between them) race on it.
static __device__ __align__(256) double g_noalloc_sums[2 * NCONV_B200_NOALLOC_MAX_C];

static int noalloc_sums_ptr(double** out) {
  static std::mutex mu;
  static double* ptr[NCONV_MAX_DEVICES] = {};
  int dev = 0;
  NCONV_RETURN_IF_ERR(cudaGetDevice(&dev));
  if (dev < 0 || dev >= NCONV_MAX_DEVICES) return NCONV_ERR_UNSUPPORTED;
  std::lock_guard<std::mutex> lock(mu);
  if (!ptr[dev]) {
    double* p = nullptr;
    NCONV_RETURN_IF_ERR(cudaGetSymbolAddress((void**)&p, g_noalloc_sums));     // the current device's copy of the array
    ptr[dev] = p;
  }
  *out = ptr[dev];
  return 0;
}

extern "C" int nconv_b200_relu_bias_bwd_noalloc(const float* gy, const float* y, float* g, float* gbias, int B, int C, int H,
                                                int W, int relu, void* stream) {
  if (!gy || (relu && !y)) return NCONV_ERR_BAD_ARG;
  if (B <= 0 || C <= 0 || H <= 0 || W <= 0) return NCONV_ERR_BAD_ARG;
  if (C > NCONV_B200_NOALLOC_MAX_C) return NCONV_ERR_STATIC_SCRATCH;           // beyond the static array: use the workspace entry point
  if ((long long)B > 65535) return NCONV_ERR_UNSUPPORTED;                       // gridDim.z
  double* sums = nullptr;
  { const int rc = noalloc_sums_ptr(&sums); if (rc) return rc; }
  cudaStream_t st = (cudaStream_t)stream;
  const int HW = H * W;
  const long long total = (long long)B * C * HW;
  NCONV_RETURN_IF_ERR(cudaMemsetAsync(sums, 0, sizeof(double) * 2 * C, st));
  {
    LaunchScope ls(st, "relu_bias_bwd (static scratch)", 2.0 * total, 4.0 * total * (g ? 3.0 : (relu ? 2.0 : 1.0)));
    relu_bias_bwd_kernel<<<dim3(red_chunks(HW, C, B), C, B), RED_THREADS, 0, st>>>(gy, y, g, sums, C, HW, relu);
  }
  if (gbias) {
    LaunchScope ls(st, "channel_sum_out", 0.0, 0.0);
    double_to_float_kernel<<<(C + 127) / 128, 128, 0, st>>>(sums, gbias, C, 1);
  }
  return (int)cudaGetLastError();
}

// =====================================================================================================
// Weight gradient:  gW[co][ci][u][v] = sum_{b,oy,ox} g[b,co,oy,ox] * x[b,ci, oy*S-P+u, ox*S-P+v]
// (Conv2d K in {1,3}, S in {1,2}; ConvTranspose2d(4,2,1) is the same sum with the roles of input and
// output gradient swapped, K=4 S=2 P=1.)
//
// A GEMM C[co][(ci,u,v)] = G[co][px] * X[px][(ci,u,v)] with the pixel axis as the contraction.  A CTA
// owns a CO_B x 8(ci) x K*K block of C and a strided subset of pixel chunks (split-K); a chunk is 64
// output pixels of one row: G tile [CO_B][64] and the x halo rows [8][K][(64-1)*S+K] go through shared
// memory, a thread accumulates 4(co) x 1(ci) x K*K taps in registers over its 64/NSUB pixels with the
// x window sliding in registers (S new columns per pixel), then the NSUB pixel sub-ranges are summed
// through shared memory and the CTA writes its partial block; wgrad_reduce sums the split-K partials
// in a fixed order (deterministic).
// =====================================================================================================
namespace nconv {

constexpr int WG_CH = 64;        // output pixels per chunk
constexpr int WG_CIB = 8;        // input channels per CTA

struct WgradArgs {
  const float* g; const float* x; float* part;
  int B, CI, CO, H, W, HO, WO, nsplit, chunks_x;
  long long chunks;
};

template <int K, int S, int TMG>
struct WgradCfg {
  static constexpr int CO_B = 4 * TMG, NSUB = 256 / (TMG * WG_CIB), PXT = WG_CH / NSUB;
  static constexpr int XC = (WG_CH - 1) * S + K;            // staged x columns
  static constexpr int XCP = XC | 1;                        // odd row pitch: the 8 channels of a warp hit 8 distinct banks
  static constexpr int GP = WG_CH + 1;
  static constexpr int PER = 4 * K * K;
  static constexpr int TILE = CO_B * GP + WG_CIB * K * XCP, RED = TMG * WG_CIB * PER;
  static constexpr int NBUF = 2;                            // double-buffered tiles (dynamic shared memory)
  static constexpr size_t SMEM = sizeof(float) * ((NBUF * TILE > RED) ? NBUF * TILE : RED);
};

template <int K, int S, int P, int TMG>
__global__ void __launch_bounds__(256) conv_wgrad_kernel(const WgradArgs a) {
  using CFG = WgradCfg<K, S, TMG>;
  constexpr int CO_B = CFG::CO_B, NSUB = CFG::NSUB, PXT = CFG::PXT, XC = CFG::XC, XCP = CFG::XCP, GP = CFG::GP;
  constexpr int TILE = CFG::TILE;
  static_assert(NSUB >= 1 && PXT * NSUB == WG_CH, "bad wgrad tiling");
  // Double-buffered tiles: the copies of the CTA's next chunk are issued with cp.async before the current chunk is
  // accumulated.  (ncu before: 42 % of the warp samples were STS instructions waiting for their global load — the
  // staging loops were load -> store per element — and FFMA2 only 23 %.)
  constexpr bool DB = true;
  extern __shared__ __align__(16) float s_all[];

  const int tid = threadIdx.x;
  const int ci_l = tid % WG_CIB, tm = (tid / WG_CIB) % TMG, sub = tid / (WG_CIB * TMG);
  const int ci0 = blockIdx.y * WG_CIB, co0 = blockIdx.z * CO_B;

  // Two-level accumulation: `acc` sums one chunk (PXT products), `tot` sums the chunks.  The rounding
  // error then grows with sqrt(PXT) + sqrt(chunks per CTA) instead of sqrt(all pixels of the CTA): these
  // sums run over up to 10^5 products of un-normalised activations with heavy cancellation.
  // The chunk sums run packed (FFMA2, common.cuh): the gradients of an output-channel PAIR are one 64-bit word
  // in shared memory, the activation is the scalar-broadcast operand.
  float acc[4][K * K], tot[4][K * K];
  u64 acc2[2][K * K];
#pragma unroll
  for (int j = 0; j < 4; ++j)
#pragma unroll
    for (int t = 0; t < K * K; ++t) { acc[j][t] = 0.f; tot[j][t] = 0.f; }
#pragma unroll
  for (int j = 0; j < 2; ++j)
#pragma unroll
    for (int t = 0; t < K * K; ++t) acc2[j][t] = 0ull;

  // stage chunk `ch` into `tile`: ASYNC = cp.async (4-byte copies, zero-filled outside), else load + store
  auto stage = [&](long long ch, float* tile, auto async_tag) {
    constexpr bool ASYNC = decltype(async_tag)::value;
    float* t_g = tile;
    float* t_x = tile + CO_B * GP;
    const int cx = (int)(ch % a.chunks_x);
    const long long r = ch / a.chunks_x;
    const int oy = (int)(r % a.HO), b = (int)(r / a.HO);
    const int ox0 = cx * WG_CH;
    // g tile: a thread keeps its pixel column and walks the output channels 4 apart (256 threads = 4 rows of 64 pixels):
    // source and destination advance by constants, no division per element
    {
      const int px = tid % WG_CH, cb = tid / WG_CH;                 // cb in 0..3
      const bool col_ok = ox0 + px < a.WO;
      const size_t plane = (size_t)a.HO * a.WO;
      const float* src = a.g + (((size_t)b * a.CO + co0 + cb) * a.HO + oy) * a.WO + ox0 + px;
      float* dst = t_g + (cb >> 1) * (2 * GP) + 2 * px + (cb & 1);   // [co/2][px][co&1]
#pragma unroll
      for (int k = 0; k < CO_B / 4; ++k) {
        const bool ok = col_ok && co0 + cb + 4 * k < a.CO;
        if constexpr (ASYNC) cp_async4(dst, ok ? src : a.g, ok); else *dst = ok ? __ldg(src) : 0.f;
        src += 4 * plane;
        dst += 2 * (2 * GP);
      }
    }
    // x halo rows: warp w takes rows w, w+8, ... of the WG_CIB*K staged rows, lanes walk the columns
    {
      const int lane = tid & 31, wrp = tid >> 5;
#pragma unroll
      for (int i = 0; i < (WG_CIB * K + 7) / 8; ++i) {
        const int row = wrp + 8 * i;
        if (row >= WG_CIB * K) break;
        const int cc = row / K, u = row - cc * K;
        const int iy = oy * S - P + u;
        const bool row_ok = ci0 + cc < a.CI && iy >= 0 && iy < a.H;
        const float* srow = a.x + (row_ok ? (((size_t)b * a.CI + ci0 + cc) * a.H + iy) * a.W : 0);
        float* drow = t_x + row * XCP;
        const int ix0 = ox0 * S - P;
#pragma unroll
        for (int c = 0; c < (XC + 31) / 32; ++c) {
          const int xc = lane + 32 * c;
          if (xc < XC) {
            const int ix = ix0 + xc;
            const bool ok = row_ok && ix >= 0 && ix < a.W;
            if constexpr (ASYNC) cp_async4(drow + xc, ok ? srow + ix : a.x, ok); else drow[xc] = ok ? __ldg(srow + ix) : 0.f;
          }
        }
      }
    }
  };

  if constexpr (DB) {
    if ((long long)blockIdx.x < a.chunks) stage(blockIdx.x, s_all, std::true_type{});
  }
  int it = 0;
  for (long long ch = blockIdx.x; ch < a.chunks; ch += a.nsplit, ++it) {
    float* tile = s_all;
    if constexpr (DB) {
      tile = s_all + (it & 1) * TILE;
      cp_async_wait_all();
      __syncthreads();            // chunk `ch` visible to all; everyone has finished the previous chunk (other buffer free)
      if (ch + a.nsplit < a.chunks) stage(ch + a.nsplit, s_all + ((it & 1) ^ 1) * TILE, std::true_type{});
    } else {
      __syncthreads();
      stage(ch, tile, std::false_type{});
      __syncthreads();
    }
    const float* s_g = tile;
    const float* s_x = tile + CO_B * GP;
    const float* gx = s_g + (2 * tm) * (2 * GP) + 2 * sub * PXT;     // pairs (4tm, 4tm+1), (4tm+2, 4tm+3)
    const float* xx = s_x + ci_l * K * XCP + sub * PXT * S;
    float win[K][K];
#pragma unroll
    for (int u = 0; u < K; ++u)
#pragma unroll
      for (int v = 0; v < K - S; ++v) win[u][v + S] = xx[u * XCP + v];      // pre-shifted: first step shifts left by S
#pragma unroll
    for (int p = 0; p < PXT; ++p) {
#pragma unroll
      for (int u = 0; u < K; ++u) {
#pragma unroll
        for (int v = 0; v < K - S; ++v) win[u][v] = win[u][v + S];
#pragma unroll
        for (int v = (K - S > 0 ? K - S : 0); v < K; ++v) win[u][v] = xx[u * XCP + p * S + v];
      }
      u64 gv2[2];
#pragma unroll
      for (int j = 0; j < 2; ++j) gv2[j] = *reinterpret_cast<const u64*>(gx + j * (2 * GP) + 2 * p);
#pragma unroll
      for (int j = 0; j < 2; ++j)
#pragma unroll
        for (int u = 0; u < K; ++u)
#pragma unroll
          for (int v = 0; v < K; ++v) ffma2(acc2[j][u * K + v], gv2[j], win[u][v]);
    }
#pragma unroll
    for (int j = 0; j < 2; ++j)
#pragma unroll
      for (int t = 0; t < K * K; ++t) {
        tot[2 * j][t] += lo32f(acc2[j][t]); tot[2 * j + 1][t] += hi32f(acc2[j][t]);
        acc2[j][t] = 0ull;
      }
  }
#pragma unroll
  for (int j = 0; j < 4; ++j)
#pragma unroll
    for (int t = 0; t < K * K; ++t) acc[j][t] = tot[j][t];

  // ---- sum the NSUB pixel sub-ranges, write the CTA's partial block
  __syncthreads();
  float* s_red = s_all;    // one sub-range at a time: TMG*8 * 4*K*K floats
  for (int s = NSUB - 1; s >= 1; --s) {
    if (sub == s) {
#pragma unroll
      for (int j = 0; j < 4; ++j)
#pragma unroll
        for (int t = 0; t < K * K; ++t) s_red[(j * K * K + t) * (TMG * WG_CIB) + tm * WG_CIB + ci_l] = acc[j][t];
    }
    __syncthreads();
    if (sub == 0) {
#pragma unroll
      for (int j = 0; j < 4; ++j)
#pragma unroll
        for (int t = 0; t < K * K; ++t) acc[j][t] += s_red[(j * K * K + t) * (TMG * WG_CIB) + tm * WG_CIB + ci_l];
    }
    __syncthreads();
  }
  if (sub == 0 && ci0 + ci_l < a.CI) {
    float* out = a.part + (size_t)blockIdx.x * a.CO * a.CI * K * K;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int co = co0 + 4 * tm + j;
      if (co < a.CO) {
#pragma unroll
        for (int t = 0; t < K * K; ++t) out[((size_t)co * a.CI + ci0 + ci_l) * K * K + t] = acc[j][t];
      }
    }
  }
}

__global__ void wgrad_reduce_kernel(const float* __restrict__ part, float* __restrict__ gw, int n, int nsplit) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double s = 0.0;
  for (int k = 0; k < nsplit; ++k) s += (double)part[(size_t)k * n + i];
  gw[i] = (float)s;
}

static inline int wgrad_nsplit(long long chunks, int ciblocks, int coblocks) {
  long long want = (148 * 4 + (long long)ciblocks * coblocks - 1) / ((long long)ciblocks * coblocks);
  if (want > chunks) want = chunks;
  if (want > 512) want = 512;
  return (int)(want < 1 ? 1 : want);
}

// CO == 1, 3x3 / stride 1 / pad 1 (the C->1 depth heads, step2.py:156): the GEMM form above would spend 31 of its 32
// output-channel rows on padding (measured 0.56 TFLOP/s, 3.2 ms per launch at 32->1 480x640).  Here the sum is
// streamed: gw[ci][u][v] = sum_p x[ci][p] * g[p - (u-1, v-1)] — a thread takes one pixel p, loads the 3x3
// neighbourhood of the single-channel gradient once and 8 input channels of x (coalesced), 72 FMAs; CTAs stride over
// image rows (split-K), block-reduce their 8x9 sums and wgrad_reduce_kernel adds the partials in a fixed order.
constexpr int W1_CIG = 8;
__global__ void __launch_bounds__(256) conv_wgrad_co1_kernel(const float* __restrict__ g, const float* __restrict__ x,
                                                             float* __restrict__ part, int B, int CI, int H, int W) {
  __shared__ float s_red[8][W1_CIG * 9];
  const int ci0 = blockIdx.y * W1_CIG;
  float acc[W1_CIG][9];
#pragma unroll
  for (int j = 0; j < W1_CIG; ++j)
#pragma unroll
    for (int t = 0; t < 9; ++t) acc[j][t] = 0.f;
  for (int unit = blockIdx.x; unit < B * H; unit += gridDim.x) {
    const int b = unit / H, y = unit - b * H;
    const float* gb = g + (size_t)b * H * W;
    for (int xx = threadIdx.x; xx < W; xx += 256) {
      float gn[9];
#pragma unroll
      for (int u = 0; u < 3; ++u)
#pragma unroll
        for (int v = 0; v < 3; ++v) {
          const int oy = y - (u - 1), ox = xx - (v - 1);
          gn[u * 3 + v] = (oy >= 0 && oy < H && ox >= 0 && ox < W) ? __ldg(gb + (size_t)oy * W + ox) : 0.f;
        }
      float xv[W1_CIG];
#pragma unroll
      for (int j = 0; j < W1_CIG; ++j)
        xv[j] = (ci0 + j < CI) ? __ldg(x + (((size_t)b * CI + ci0 + j) * H + y) * W + xx) : 0.f;
#pragma unroll
      for (int j = 0; j < W1_CIG; ++j)
#pragma unroll
        for (int t = 0; t < 9; ++t) acc[j][t] = fmaf(xv[j], gn[t], acc[j][t]);
    }
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
  for (int j = 0; j < W1_CIG; ++j)
#pragma unroll
    for (int t = 0; t < 9; ++t) {
      const float v = warp_sum(acc[j][t]);
      if (lane == 0) s_red[warp][j * 9 + t] = v;
    }
  __syncthreads();
  if (threadIdx.x < W1_CIG * 9) {
    float v = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) v += s_red[k][threadIdx.x];
    const int j = threadIdx.x / 9, t = threadIdx.x - j * 9;
    if (ci0 + j < CI) part[(size_t)blockIdx.x * CI * 9 + (size_t)(ci0 + j) * 9 + t] = v;
  }
}

static int launch_wgrad_co1(const float* g, const float* x, float* gw, void* ws, size_t ws_bytes, int B, int CI, int H,
                            int W, cudaStream_t st) {
  const int ciblocks = ceil_div(CI, W1_CIG);
  int nsplit = (148 * 4 + ciblocks - 1) / ciblocks;
  if (nsplit > B * H) nsplit = B * H;
  if (nsplit > 512) nsplit = 512;
  const size_t n = (size_t)CI * 9;
  if (ws_bytes < n * nsplit * sizeof(float)) return NCONV_ERR_WORKSPACE;
  char label[64];
  snprintf(label, sizeof(label), "conv_wgrad k3 s1 %d->1 @%dx%d", CI, H, W);
  {
    LaunchScope ls(st, label, 2.0 * CI * 9 * (double)B * H * W, 4.0 * ((double)B * CI * H * W + (double)B * H * W));
    conv_wgrad_co1_kernel<<<dim3(nsplit, ciblocks), 256, 0, st>>>(g, x, (float*)ws, B, CI, H, W);
  }
  {
    LaunchScope ls(st, "wgrad_reduce", (double)n * nsplit, 4.0 * n * (nsplit + 1));
    wgrad_reduce_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>((const float*)ws, gw, (int)n, nsplit);
  }
  return (int)cudaGetLastError();
}

template <int K, int S, int P>
static int launch_wgrad(const float* g, const float* x, float* gw, void* ws, size_t ws_bytes, int B, int CI, int CO, int H,
                        int W, int HO, int WO, cudaStream_t st) {
  WgradArgs a;
  a.g = g; a.x = x; a.part = (float*)ws; a.B = B; a.CI = CI; a.CO = CO; a.H = H; a.W = W; a.HO = HO; a.WO = WO;
  a.chunks_x = ceil_div(WO, WG_CH);
  a.chunks = (long long)B * HO * a.chunks_x;
  const bool small_co = CO <= 32;
  const int co_b = small_co ? 32 : 64;
  const int ciblocks = ceil_div(CI, WG_CIB), coblocks = ceil_div(CO, co_b);
  a.nsplit = wgrad_nsplit(a.chunks, ciblocks, coblocks);
  const size_t n = (size_t)CO * CI * K * K;
  if (ws_bytes < n * a.nsplit * sizeof(float)) return NCONV_ERR_WORKSPACE;
  {
    static bool attr_done[NCONV_MAX_DEVICES] = {};
    int dev = 0;
    NCONV_RETURN_IF_ERR(cudaGetDevice(&dev));
    if (dev >= NCONV_MAX_DEVICES || !attr_done[dev]) {
      NCONV_RETURN_IF_ERR(cudaFuncSetAttribute(conv_wgrad_kernel<K, S, P, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                               (int)WgradCfg<K, S, 8>::SMEM));
      NCONV_RETURN_IF_ERR(cudaFuncSetAttribute(conv_wgrad_kernel<K, S, P, 16>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                               (int)WgradCfg<K, S, 16>::SMEM));
      if (dev < NCONV_MAX_DEVICES) attr_done[dev] = true;
    }
  }
  char label[64];
  snprintf(label, sizeof(label), "conv_wgrad k%d s%d %d->%d @%dx%d", K, S, CI, CO, H, W);
  {
    LaunchScope ls(st, label, 2.0 * CI * CO * K * K * (double)B * HO * WO,
                   4.0 * ((double)B * CI * H * W + (double)B * CO * HO * WO));
    dim3 grid(a.nsplit, ciblocks, coblocks);
    if (small_co) conv_wgrad_kernel<K, S, P, 8><<<grid, 256, WgradCfg<K, S, 8>::SMEM, st>>>(a);
    else conv_wgrad_kernel<K, S, P, 16><<<grid, 256, WgradCfg<K, S, 16>::SMEM, st>>>(a);
  }
  {
    LaunchScope ls(st, "wgrad_reduce", (double)n * a.nsplit, 4.0 * n * (a.nsplit + 1));
    wgrad_reduce_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>((const float*)ws, gw, (int)n, a.nsplit);
  }
  return (int)cudaGetLastError();
}

}  // namespace nconv

// Workspace for nconv_b200_conv_wgrad (split-K partials).
extern "C" size_t nconv_b200_conv_wgrad_ws_bytes(int CI, int CO, int K) { return (size_t)512 * CO * CI * K * K * sizeof(float); }

// gw [CO][CI][K][K] = sum g[b,co,oy,ox] * x[b,ci,oy*S-P+u,ox*S-P+v];  g: [B,CO,HO,WO], x: [B,CI,H,W].
// (K,S,P) in {(1,1,0),(1,2,0),(3,1,1),(3,2,1),(4,2,1)}.
extern "C" int nconv_b200_conv_wgrad(const float* g, const float* x, float* gw, void* ws, size_t ws_bytes, int B, int CI,
                                     int CO, int H, int W, int K, int S, int P, void* stream) {
  if (!g || !x || !gw || !ws || B <= 0 || CI <= 0 || CO <= 0 || H <= 0 || W <= 0) return NCONV_ERR_BAD_ARG;
  const int HO = (H + 2 * P - K) / S + 1, WO = (W + 2 * P - K) / S + 1;
  if (HO <= 0 || WO <= 0) return NCONV_ERR_BAD_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  if (K == 3 && S == 1 && P == 1 && CO == 1) return nconv::launch_wgrad_co1(g, x, gw, ws, ws_bytes, B, CI, H, W, st);
  if (K == 3 && S == 1 && P == 1) return nconv::launch_wgrad<3, 1, 1>(g, x, gw, ws, ws_bytes, B, CI, CO, H, W, HO, WO, st);
  if (K == 3 && S == 2 && P == 1) return nconv::launch_wgrad<3, 2, 1>(g, x, gw, ws, ws_bytes, B, CI, CO, H, W, HO, WO, st);
  if (K == 1 && S == 1 && P == 0) return nconv::launch_wgrad<1, 1, 0>(g, x, gw, ws, ws_bytes, B, CI, CO, H, W, HO, WO, st);
  if (K == 1 && S == 2 && P == 0) return nconv::launch_wgrad<1, 2, 0>(g, x, gw, ws, ws_bytes, B, CI, CO, H, W, HO, WO, st);
  if (K == 4 && S == 2 && P == 1) return nconv::launch_wgrad<4, 2, 1>(g, x, gw, ws, ws_bytes, B, CI, CO, H, W, HO, WO, st);
  return NCONV_ERR_UNSUPPORTED;
}

// =====================================================================================================
// Gather convolutions for the data gradients that are not a stride-1 "same" convolution:
//   MODE 0 (transposed gather): dgrad of Conv2d(K, stride S, pad P):
//       out[b,oc,iy,ix] = sum_{rc,u,v} g[b,rc,(iy+P-u)/S,(ix+P-v)/S] * W[rc][oc][u][v]   (when divisible, in range)
//     one CTA works on ONE output phase (iy%S, ix%S), so the set of contributing taps is uniform and
//     the gradient reads are unit-stride;
//   MODE 1 (strided gather): dgrad of ConvTranspose2d(K=4, S=2, P=1), a strided convolution of g:
//       out[b,oc,iy,ix] = sum_{rc,kh,kw} g[b,rc,iy*S-P+kh,ix*S-P+kw] * Wt[oc][rc][kh][kw]
// rc = channels of g (the layer's outputs), oc = channels of the result (the layer's inputs).
// Weights are re-packed once per call to wt[tap][rc][oc] (gather_pack_kernel); a thread owns 4 pixels x
// 8 oc, reduction channels go through shared memory 8 at a time, g is read through L1.
// =====================================================================================================
namespace nconv {

struct GatherArgs {
  const float* g; const float* wt; float* out;
  int B, RC, OC, GH, GW, OH, OW, K, S, P;
};

// conv weight W[co][ci][K][K] (transposed=0) or ConvTranspose weight Wt[ci][co][K][K] (transposed=1)
// -> wt[tap][rc=co][oc=ci]
__global__ void gather_pack_kernel(const float* __restrict__ w, float* __restrict__ wt, int CO, int CI, int KK, int transposed) {
  const int total = KK * CO * CI;
  for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += gridDim.x * blockDim.x) {
    const int ci = idx % CI, co = (idx / CI) % CO, tap = idx / (CI * CO);
    wt[idx] = transposed ? __ldg(w + ((size_t)ci * CO + co) * KK + tap) : __ldg(w + ((size_t)co * CI + ci) * KK + tap);
  }
}

__host__ __device__ __forceinline__ int floordiv(int a, int b) { return (a >= 0) ? a / b : -((-a + b - 1) / b); }

// KT/ST/PT > 0 fix (K, S, P) at compile time (the shapes of the RGB-guided stack: 3/2/1, 1/2/0, 4/2/1): the tap loops
// unroll, the phase tests become bit operations and the loads of a reduction channel's taps are issued together.  The
// generic form (0/0/0, runtime K/S/P) spent its issue slots on integer division (ncu: IMAD/BRA/ISETP 60 % of the warp
// samples, FFMA 10 %).
template <int MODE, int KT, int ST, int PT>
__global__ void __launch_bounds__(256) conv_gather_kernel(const GatherArgs a) {
  __shared__ __align__(16) float s_w[16 * 8 * 8];      // [tap][rc 8][oc 8]
  const int K = KT ? KT : a.K, S = ST ? ST : a.S, P = KT ? PT : a.P, KK = K * K;
  const int phases = (MODE == 0) ? S * S : 1;
  const int nob = (a.OC + 7) / 8;
  const int z = blockIdx.z;
  const int ph = z % phases, ob = (z / phases) % nob, b = z / (phases * nob);
  const int py = ph / S, px = ph % S;
  const int PH = (MODE == 0) ? (a.OH - py + S - 1) / S : a.OH;
  const int PW = (MODE == 0) ? (a.OW - px + S - 1) / S : a.OW;
  const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
  const int q4 = 4 * (blockIdx.x * 32 + lane), j = blockIdx.y * 8 + wrp;
  const bool active = (j < PH) && (q4 < PW);

  float acc[8][4];
#pragma unroll
  for (int o = 0; o < 8; ++o)
#pragma unroll
    for (int p = 0; p < 4; ++p) acc[o][p] = 0.f;

  for (int rc0 = 0; rc0 < a.RC; rc0 += 8) {
    __syncthreads();
    for (int idx = threadIdx.x; idx < KK * 64; idx += 256) {
      const int o = idx & 7, r = (idx >> 3) & 7, tap = idx >> 6;
      float v = 0.f;
      if (rc0 + r < a.RC && ob * 8 + o < a.OC) v = __ldg(a.wt + ((size_t)tap * a.RC + rc0 + r) * a.OC + ob * 8 + o);
      s_w[idx] = v;
    }
    __syncthreads();
    if (!active) continue;
    const int nr = min(8, a.RC - rc0);
    for (int r = 0; r < nr; ++r) {
      const float* gp = a.g + ((size_t)b * a.RC + rc0 + r) * a.GH * a.GW;
#pragma unroll
      for (int u = 0; u < K; ++u) {
        int gy;
        if (MODE == 0) {
          const int t = py + P - u;
          if (((t % S) + S) % S != 0) continue;
          gy = j + floordiv(t, S);
        } else {
          gy = j * S - P + u;
        }
        if (gy < 0 || gy >= a.GH) continue;
        const float* grow = gp + (size_t)gy * a.GW;
#pragma unroll
        for (int v = 0; v < K; ++v) {
          int gx0, step;
          if (MODE == 0) {
            const int t = px + P - v;
            if (((t % S) + S) % S != 0) continue;
            gx0 = q4 + floordiv(t, S); step = 1;
          } else {
            gx0 = q4 * S - P + v; step = S;
          }
          float gv[4];
#pragma unroll
          for (int p = 0; p < 4; ++p) {
            const int gx = gx0 + p * step;
            gv[p] = (gx >= 0 && gx < a.GW) ? __ldg(grow + gx) : 0.f;
          }
          const float4 w0 = *reinterpret_cast<const float4*>(s_w + ((u * K + v) * 8 + r) * 8);
          const float4 w1 = *reinterpret_cast<const float4*>(s_w + ((u * K + v) * 8 + r) * 8 + 4);
          const float wv[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
          for (int o = 0; o < 8; ++o)
#pragma unroll
            for (int p = 0; p < 4; ++p) acc[o][p] = fmaf(wv[o], gv[p], acc[o][p]);
        }
      }
    }
  }
  if (!active) return;
  const int iy = (MODE == 0) ? j * S + py : j;
#pragma unroll
  for (int o = 0; o < 8; ++o) {
    const int oc = ob * 8 + o;
    if (oc >= a.OC) break;
    float* orow = a.out + (((size_t)b * a.OC + oc) * a.OH + iy) * a.OW;
#pragma unroll
    for (int p = 0; p < 4; ++p) {
      if (q4 + p < PW) orow[(MODE == 0) ? (q4 + p) * S + px : q4 + p] = acc[o][p];
    }
  }
}

// Data gradient of ConvTranspose2d(4, 2, 1) — the strided 4x4 gather of MODE 1 — specialised: per reduction channel
// and kernel row a thread reads the 10 CONTIGUOUS gradient columns its 4 output pixels need (2 x LDG.128 + 2 scalars;
// the generic kernel issued 16 stride-2 scalar loads for the same 128 products) and runs the products packed over
// output-channel pairs (FFMA2).  Same thread/CTA geometry and weight chunking as conv_gather_kernel<1, ...>.
__global__ void __launch_bounds__(256) convT_dgrad_k4s2_kernel(const GatherArgs a) {
  __shared__ __align__(16) float s_w[16 * 8 * 8];      // [tap][rc 8][oc 8]
  const int nob = (a.OC + 7) / 8;
  const int z = blockIdx.z;
  const int ob = z % nob, b = z / nob;
  const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
  const int q4 = 4 * (blockIdx.x * 32 + lane), j = blockIdx.y * 8 + wrp;
  const bool active = (j < a.OH) && (q4 < a.OW);
  const bool vec = (a.GW % 4 == 0) && ((reinterpret_cast<uintptr_t>(a.g) & 15) == 0) && (2 * q4 + 7 < a.GW);

  u64 acc2[4][4];                                      // [oc pair][pixel]
#pragma unroll
  for (int o = 0; o < 4; ++o)
#pragma unroll
    for (int p = 0; p < 4; ++p) acc2[o][p] = 0ull;

  for (int rc0 = 0; rc0 < a.RC; rc0 += 8) {
    __syncthreads();
    for (int idx = threadIdx.x; idx < 16 * 64; idx += 256) {
      const int o = idx & 7, r = (idx >> 3) & 7, tap = idx >> 6;
      float v = 0.f;
      if (rc0 + r < a.RC && ob * 8 + o < a.OC) v = __ldg(a.wt + ((size_t)tap * a.RC + rc0 + r) * a.OC + ob * 8 + o);
      s_w[idx] = v;
    }
    __syncthreads();
    if (!active) continue;
    const int nr = min(8, a.RC - rc0);
    for (int r = 0; r < nr; ++r) {
      const float* gp = a.g + ((size_t)b * a.RC + rc0 + r) * a.GH * a.GW;
#pragma unroll
      for (int kh = 0; kh < 4; ++kh) {
        const int gy = 2 * j - 1 + kh;
        if (gy < 0 || gy >= a.GH) continue;
        const float* grow = gp + (size_t)gy * a.GW;
        const int c0 = 2 * q4 - 1;                     // column of gv[0]
        float gv[10];
        if (vec) {
          const float4 m0 = __ldg(reinterpret_cast<const float4*>(grow + c0 + 1));
          const float4 m1 = __ldg(reinterpret_cast<const float4*>(grow + c0 + 5));
          gv[0] = (c0 >= 0) ? __ldg(grow + c0) : 0.f;
          gv[1] = m0.x; gv[2] = m0.y; gv[3] = m0.z; gv[4] = m0.w;
          gv[5] = m1.x; gv[6] = m1.y; gv[7] = m1.z; gv[8] = m1.w;
          gv[9] = (c0 + 9 < a.GW) ? __ldg(grow + c0 + 9) : 0.f;
        } else {
#pragma unroll
          for (int t = 0; t < 10; ++t) gv[t] = (c0 + t >= 0 && c0 + t < a.GW) ? __ldg(grow + c0 + t) : 0.f;
        }
#pragma unroll
        for (int kw = 0; kw < 4; ++kw) {
          const ulonglong2 w01 = *reinterpret_cast<const ulonglong2*>(s_w + ((kh * 4 + kw) * 8 + r) * 8);
          const ulonglong2 w23 = *reinterpret_cast<const ulonglong2*>(s_w + ((kh * 4 + kw) * 8 + r) * 8 + 4);
#pragma unroll
          for (int p = 0; p < 4; ++p) {
            const float x = gv[2 * p + kw];
            ffma2(acc2[0][p], w01.x, x); ffma2(acc2[1][p], w01.y, x);
            ffma2(acc2[2][p], w23.x, x); ffma2(acc2[3][p], w23.y, x);
          }
        }
      }
    }
  }
  if (!active) return;
#pragma unroll
  for (int o = 0; o < 8; ++o) {
    const int oc = ob * 8 + o;
    if (oc >= a.OC) break;
    float* orow = a.out + (((size_t)b * a.OC + oc) * a.OH + j) * a.OW;
#pragma unroll
    for (int p = 0; p < 4; ++p)
      if (q4 + p < a.OW) orow[q4 + p] = (o & 1) ? hi32f(acc2[o >> 1][p]) : lo32f(acc2[o >> 1][p]);
  }
}

// Data gradient of Conv2d(3, 2, 1) specialised.  conv_gather_kernel<0,3,2,1> runs one CTA per output phase: the four
// phases carry 1/2/2/4 taps (unbalanced CTAs), and each writes every second column (half-used sectors).  Here a thread
// owns a gradient row j and 2 gradient columns q2, q2+1 and produces ALL four phases under them — output rows 2j, 2j+1 x
// columns 2q2 .. 2q2+3 x 8 oc — from the 2 x 3 gradient values g[j..j+1][q2..q2+2] (LDG.64 + LDG.32 per row, prefetched
// one reduction channel ahead); the 9 taps run as FFMA2 over oc pairs and the results leave as one STG.128 per row.
//   row 2j   <- (u=1, g row j);   row 2j+1   <- (u=2, g row j), (u=0, g row j+1);   columns alike with v / q.
__global__ void __launch_bounds__(256) conv_dgrad_k3s2_kernel(const GatherArgs a) {
  __shared__ __align__(16) float s_w[9 * 8 * 8];       // [tap][rc 8][oc 8]
  const int nob = (a.OC + 7) / 8;
  const int z = blockIdx.z;
  const int ob = z % nob, b = z / nob;
  const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
  const int q2 = 2 * (blockIdx.x * 32 + lane), j = blockIdx.y * 8 + wrp;
  const bool active = (j < a.GH) && (2 * j < a.OH) && (q2 < a.GW) && (2 * q2 < a.OW);
  const bool vec2 = (a.GW % 2 == 0) && ((reinterpret_cast<uintptr_t>(a.g) & 7) == 0) && (q2 + 1 < a.GW);
  const bool row1 = (j + 1 < a.GH);
  const bool col2 = (q2 + 2 < a.GW);

  u64 acc2[2][4][4];                                   // [output row][output column][oc pair]
#pragma unroll
  for (int y = 0; y < 2; ++y)
#pragma unroll
    for (int x = 0; x < 4; ++x)
#pragma unroll
      for (int o = 0; o < 4; ++o) acc2[y][x][o] = 0ull;

  auto load6 = [&](const float* gp, float (&G)[2][3]) {
    const float* r0 = gp + (size_t)j * a.GW + q2;
    if (vec2) {
      const float2 m = __ldg(reinterpret_cast<const float2*>(r0));
      G[0][0] = m.x; G[0][1] = m.y;
    } else {
      G[0][0] = __ldg(r0); G[0][1] = (q2 + 1 < a.GW) ? __ldg(r0 + 1) : 0.f;
    }
    G[0][2] = col2 ? __ldg(r0 + 2) : 0.f;
    if (row1) {
      const float* r1 = r0 + a.GW;
      if (vec2) {
        const float2 m = __ldg(reinterpret_cast<const float2*>(r1));
        G[1][0] = m.x; G[1][1] = m.y;
      } else {
        G[1][0] = __ldg(r1); G[1][1] = (q2 + 1 < a.GW) ? __ldg(r1 + 1) : 0.f;
      }
      G[1][2] = col2 ? __ldg(r1 + 2) : 0.f;
    } else {
      G[1][0] = G[1][1] = G[1][2] = 0.f;
    }
  };

  const size_t plane = (size_t)a.GH * a.GW;
  for (int rc0 = 0; rc0 < a.RC; rc0 += 8) {
    __syncthreads();
    for (int idx = threadIdx.x; idx < 9 * 64; idx += 256) {
      const int o = idx & 7, r = (idx >> 3) & 7, tap = idx >> 6;
      float v = 0.f;
      if (rc0 + r < a.RC && ob * 8 + o < a.OC) v = __ldg(a.wt + ((size_t)tap * a.RC + rc0 + r) * a.OC + ob * 8 + o);
      s_w[idx] = v;
    }
    __syncthreads();
    if (!active) continue;
    const int nr = min(8, a.RC - rc0);
    const float* gp = a.g + ((size_t)b * a.RC + rc0) * plane;
    float G[2][3], Gn[2][3];
    load6(gp, G);
    for (int r = 0; r < nr; ++r) {
      if (r + 1 < nr) load6(gp + (size_t)(r + 1) * plane, Gn);
#pragma unroll
      for (int u = 0; u < 3; ++u) {
        const int y = (u == 1) ? 0 : 1, gr = (u == 0) ? 1 : 0;
#pragma unroll
        for (int v = 0; v < 3; ++v) {
          const int xo = (v == 1) ? 0 : 1, gc = (v == 0) ? 1 : 0;
          const ulonglong2 w01 = *reinterpret_cast<const ulonglong2*>(s_w + ((u * 3 + v) * 8 + r) * 8);
          const ulonglong2 w23 = *reinterpret_cast<const ulonglong2*>(s_w + ((u * 3 + v) * 8 + r) * 8 + 4);
#pragma unroll
          for (int c = 0; c < 2; ++c) {
            const float x = G[gr][c + gc];
            ffma2(acc2[y][2 * c + xo][0], w01.x, x); ffma2(acc2[y][2 * c + xo][1], w01.y, x);
            ffma2(acc2[y][2 * c + xo][2], w23.x, x); ffma2(acc2[y][2 * c + xo][3], w23.y, x);
          }
        }
      }
#pragma unroll
      for (int y = 0; y < 2; ++y)
#pragma unroll
        for (int c = 0; c < 3; ++c) G[y][c] = Gn[y][c];
    }
  }
  if (!active) return;
  const int ix0 = 2 * q2;
  const bool st4 = (a.OW % 4 == 0) && ((reinterpret_cast<uintptr_t>(a.out) & 15) == 0) && (ix0 + 3 < a.OW);
#pragma unroll
  for (int o = 0; o < 8; ++o) {
    const int oc = ob * 8 + o;
    if (oc >= a.OC) break;
#pragma unroll
    for (int y = 0; y < 2; ++y) {
      const int iy = 2 * j + y;
      if (iy >= a.OH) break;
      float* orow = a.out + (((size_t)b * a.OC + oc) * a.OH + iy) * a.OW + ix0;
      float r4[4];
#pragma unroll
      for (int x = 0; x < 4; ++x) r4[x] = (o & 1) ? hi32f(acc2[y][x][o >> 1]) : lo32f(acc2[y][x][o >> 1]);
      if (st4) {
        *reinterpret_cast<float4*>(orow) = make_float4(r4[0], r4[1], r4[2], r4[3]);
      } else {
#pragma unroll
        for (int x = 0; x < 4; ++x)
          if (ix0 + x < a.OW) orow[x] = r4[x];
      }
    }
  }
}

// W[co][ci][u][v] -> W'[ci][co][K-1-u][K-1-v]
__global__ void wflip_kernel(const float* __restrict__ w, float* __restrict__ wf, int CO, int CI, int K) {
  const int KK = K * K, total = CO * CI * KK;
  for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += gridDim.x * blockDim.x) {
    const int tap = idx % KK, co = (idx / KK) % CO, ci = idx / (KK * CO);
    wf[idx] = __ldg(w + ((size_t)co * CI + ci) * KK + (KK - 1 - tap));
  }
}

}  // namespace nconv

extern "C" size_t nconv_b200_conv_dgrad_ws_bytes(int CI, int CO, int K) { return (size_t)CI * CO * K * K * sizeof(float); }

// Data gradient of Conv2d(K, S, P=K/2) for S = 2 (transposed = 0; w = [CO][CI][K][K]; g = [B,CO,HO,WO];
// gx = [B,CI,H,W]) and of ConvTranspose2d(4,2,1) (transposed = 1; w = [CI][CO][4][4]; g = [B,CO,2H,2W];
// gx = [B,CI,H,W]).  ws: nconv_b200_conv_dgrad_ws_bytes(CI, CO, K).
extern "C" int nconv_b200_conv_dgrad_gather(const float* g, const float* w, float* gx, void* ws, int B, int CI, int CO,
                                            int H, int W, int K, int S, int P, int transposed, void* stream) {
  if (!g || !w || !gx || !ws || B <= 0 || CI <= 0 || CO <= 0 || H <= 0 || W <= 0) return NCONV_ERR_BAD_ARG;
  if (K < 1 || K > 4 || S < 1 || S > 2) return NCONV_ERR_UNSUPPORTED;
  cudaStream_t st = (cudaStream_t)stream;
  const int KK = K * K;
  {
    LaunchScope ls(st, "gather_pack", 0.0, 8.0 * CI * CO * KK);
    gather_pack_kernel<<<(CI * CO * KK + 255) / 256, 256, 0, st>>>(w, (float*)ws, CO, CI, KK, transposed);
  }
  GatherArgs a;
  a.g = g; a.wt = (const float*)ws; a.out = gx; a.B = B; a.RC = CO; a.OC = CI; a.OH = H; a.OW = W; a.K = K; a.S = S; a.P = P;
  if (transposed) { a.GH = 2 * H; a.GW = 2 * W; }
  else { a.GH = (H + 2 * P - K) / S + 1; a.GW = (W + 2 * P - K) / S + 1; }
  const int phases = transposed ? 1 : S * S;
  const int PH = transposed ? H : (H + S - 1) / S, PW = transposed ? W : (W + S - 1) / S;
  dim3 grid(nconv::ceil_div(nconv::ceil_div(PW, 4), 32), nconv::ceil_div(PH, 8), B * ((CI + 7) / 8) * phases);
  char label[64];
  snprintf(label, sizeof(label), "conv_dgrad_gather%s k%d s%d %d<-%d @%dx%d", transposed ? "T" : "", K, S, CI, CO, H, W);
  {
    const double taps = transposed ? (double)KK : (double)KK / (S * S);
    LaunchScope ls(st, label, 2.0 * CI * CO * taps * (double)B * H * W,
                   4.0 * ((double)B * CO * a.GH * a.GW + (double)B * CI * H * W));
    if (transposed && K == 4 && S == 2 && P == 1) convT_dgrad_k4s2_kernel<<<grid, 256, 0, st>>>(a);
    else if (transposed) conv_gather_kernel<1, 0, 0, 0><<<grid, 256, 0, st>>>(a);
    else if (K == 3 && S == 2 && P == 1 && !getenv("NCONV_B200_DGRAD_PHASED")) {
      dim3 g2(nconv::ceil_div(nconv::ceil_div(a.GW, 2), 32), nconv::ceil_div(a.GH, 8), B * ((CI + 7) / 8));
      conv_dgrad_k3s2_kernel<<<g2, 256, 0, st>>>(a);
    }
    else if (K == 3 && S == 2 && P == 1) conv_gather_kernel<0, 3, 2, 1><<<grid, 256, 0, st>>>(a);
    else if (K == 1 && S == 2 && P == 0) conv_gather_kernel<0, 1, 2, 0><<<grid, 256, 0, st>>>(a);
    else conv_gather_kernel<0, 0, 0, 0><<<grid, 256, 0, st>>>(a);
  }
  return (int)cudaGetLastError();
}

// wf [CI][CO][K][K] = spatially flipped, channel-transposed w [CO][CI][K][K]: conv2d_fwd(g, wf) is the data
// gradient of a stride-1 "same" convolution.
extern "C" int nconv_b200_wflip(const float* w, float* wf, int CO, int CI, int K, void* stream) {
  if (!w || !wf || CO <= 0 || CI <= 0 || K <= 0) return NCONV_ERR_BAD_ARG;
  LaunchScope ls((cudaStream_t)stream, "wflip", 0.0, 8.0 * CO * CI * K * K);
  wflip_kernel<<<(CO * CI * K * K + 255) / 256, 256, 0, (cudaStream_t)stream>>>(w, wf, CO, CI, K);
  return (int)cudaGetLastError();
}
