// HBM-bound edges of the RGB-guided stack (/root/reference/models/step2.py) around the tcgen05
// convolutions: the layers with 1 or 3 input channels or 1 output channel (F/B <= 12, SURVEY App. A
// table 2) stay on CUDA cores and convert between fp32 NCHW and the P8 layout on the fly.
//
//   small_to_p8  fp32 NCHW [B,CI<=3,H,W] -> P8 [B,CO,H,W] (+ P8Q copy)
//                y = act(scale*conv3x3(x) + shift) (+ conv1x1(x)):  rgb_encoder0 (RGBEncoder :134-154,
//                3->32 with the 1x1 residual branch) and depth_conv (ConvBlock 1->C, :221)
//   head_p8      P8 [B,C,H,W] -> fp32 [B,1,H,W]:  d_out = d + conv3x3_{C->1}(f)  (Conv3x3 :156,
//                :255-257, :274-276), also emitted as a lone-channel P8 for the next level's ConvT.
#include "common.cuh"
#include "launch_count.h"
#include "nconv_b200.h"
#include "p8.cuh"
#include <cstdio>

namespace nconv {

struct SmallArgs {
  const float* x; const float* w; const float* w1; const float* scale; const float* shift;
  __nv_bfloat16* y; __nv_bfloat16* yq;
  int B, H, W, relu;
};

constexpr int SM_THREADS = 128;

// One thread = PXT pixels (128 apart along x, so every access is coalesced) x all CO channels.
template <int CI, int CO, int PXT>
__global__ void __launch_bounds__(SM_THREADS) small_to_p8_kernel(const SmallArgs a) {
  __shared__ __align__(16) float s_w[(CI * 9 + CI) * CO];   // [(ci*9+tap) | (CI*9+ci)][co]
  __shared__ float s_scale[CO], s_shift[CO];
  const int tid = threadIdx.x;
  for (int i = tid; i < CI * 9 * CO; i += SM_THREADS) {
    const int co = i % CO, r = i / CO;                      // r = ci*9 + tap
    s_w[i] = __ldg(a.w + (size_t)co * CI * 9 + r);
  }
  for (int i = tid; i < CI * CO; i += SM_THREADS) {
    const int co = i % CO, ci = i / CO;
    s_w[CI * 9 * CO + i] = a.w1 ? __ldg(a.w1 + (size_t)co * CI + ci) : 0.f;
  }
  for (int i = tid; i < CO; i += SM_THREADS) {
    s_scale[i] = a.scale ? __ldg(a.scale + i) : 1.f;
    s_shift[i] = a.shift ? __ldg(a.shift + i) : 0.f;
  }
  __syncthreads();

  const int n = blockIdx.z, y = blockIdx.y;
  const int xb = blockIdx.x * (SM_THREADS * PXT) + tid;
  float acc[PXT][CO];
#pragma unroll
  for (int p = 0; p < PXT; ++p)
#pragma unroll
    for (int o = 0; o < CO; ++o) acc[p][o] = 0.f;
  float centre[PXT][CI];

#pragma unroll
  for (int ci = 0; ci < CI; ++ci) {
    const float* plane = a.x + ((size_t)n * CI + ci) * a.H * a.W;
#pragma unroll
    for (int u = 0; u < 3; ++u) {
      const int yy = y + u - 1;
      float in[PXT][3];
#pragma unroll
      for (int p = 0; p < PXT; ++p)
#pragma unroll
        for (int v = 0; v < 3; ++v) {
          const int xx = xb + p * SM_THREADS + v - 1;
          in[p][v] = (yy >= 0 && yy < a.H && xx >= 0 && xx < a.W) ? __ldg(plane + (size_t)yy * a.W + xx) : 0.f;
        }
      if (u == 1) {
#pragma unroll
        for (int p = 0; p < PXT; ++p) centre[p][ci] = in[p][1];
      }
#pragma unroll
      for (int v = 0; v < 3; ++v) {
        const float* wr = s_w + (ci * 9 + u * 3 + v) * CO;
#pragma unroll
        for (int o = 0; o < CO; o += 4) {
          const float4 w4 = *reinterpret_cast<const float4*>(wr + o);
#pragma unroll
          for (int p = 0; p < PXT; ++p) {
            acc[p][o] = fmaf(w4.x, in[p][v], acc[p][o]);
            acc[p][o + 1] = fmaf(w4.y, in[p][v], acc[p][o + 1]);
            acc[p][o + 2] = fmaf(w4.z, in[p][v], acc[p][o + 2]);
            acc[p][o + 3] = fmaf(w4.w, in[p][v], acc[p][o + 3]);
          }
        }
      }
    }
  }

  const P8Geom g = P8Geom::make(a.B, CO, a.H, a.W);
  const P8QGeom qg = P8QGeom::make(a.B, CO, a.H, a.W);
#pragma unroll
  for (int p = 0; p < PXT; ++p) {
    const int x = xb + p * SM_THREADS;
    if (x >= a.W) continue;
#pragma unroll
    for (int ch = 0; ch < CO / 8; ++ch) {
      float f[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const int o = ch * 8 + e;
        float v = fmaf(acc[p][o], s_scale[o], s_shift[o]);
        if (a.relu) v = fmaxf(v, 0.f);
        if (a.w1) {
#pragma unroll
          for (int ci = 0; ci < CI; ++ci) v = fmaf(s_w[CI * 9 * CO + ci * CO + o], centre[p][ci], v);
        }
        f[e] = v;
      }
      uint4 hi, lo;
      split8(f, hi, lo);
      *reinterpret_cast<uint4*>(a.y + g.plane(n, ch, 0) + g.pixel(y, x)) = hi;
      *reinterpret_cast<uint4*>(a.y + g.plane(n, ch, 1) + g.pixel(y, x)) = lo;
      if (a.yq) {
        *reinterpret_cast<uint4*>(a.yq + qg.slab(n, ch, 0) + qg.pixel(y, x)) = hi;
        *reinterpret_cast<uint4*>(a.yq + qg.slab(n, ch, 1) + qg.pixel(y, x)) = lo;
      }
    }
  }
}

struct HeadArgs {
  const __nv_bfloat16* f; const float* w; const float* residual; float* y; __nv_bfloat16* y8;
  int B, H, W;
};

// One thread = one output pixel; the P8 ring supplies the zero padding (no bounds tests on loads).
template <int C>
__global__ void __launch_bounds__(128) head_p8_kernel(const HeadArgs a) {
  __shared__ __align__(16) float s_w[C * 9];    // [chunk][tap][8]
  for (int i = threadIdx.x; i < C * 9; i += 128) {
    const int e = i & 7, tap = (i >> 3) % 9, ch = i / 72;
    s_w[i] = __ldg(a.w + (size_t)(ch * 8 + e) * 9 + tap);
  }
  __syncthreads();
  const int n = blockIdx.z, y = blockIdx.y, x = blockIdx.x * 128 + threadIdx.x;
  if (x >= a.W) return;
  const P8Geom g = P8Geom::make(a.B, C, a.H, a.W);
  float acc0 = 0.f, acc1 = 0.f;
#pragma unroll 1
  for (int ch = 0; ch < C / 8; ++ch) {
    const __nv_bfloat16* ph = a.f + g.plane(n, ch, 0);
    const __nv_bfloat16* pl = a.f + g.plane(n, ch, 1);
#pragma unroll
    for (int tap = 0; tap < 9; ++tap) {
      const int u = tap / 3, v = tap - 3 * u;
      const size_t off = ((size_t)(y + u) * g.Wp + (x + v)) * 8;     // padded coords: (y+u-1)+1, (x+v-1)+1
      const uint4 hi = __ldg(reinterpret_cast<const uint4*>(ph + off));
      const uint4 lo = __ldg(reinterpret_cast<const uint4*>(pl + off));
      float f[8];
      join8(hi, lo, f);
      const float4 w0 = *reinterpret_cast<const float4*>(s_w + (ch * 9 + tap) * 8);
      const float4 w1 = *reinterpret_cast<const float4*>(s_w + (ch * 9 + tap) * 8 + 4);
      acc0 = fmaf(w0.x, f[0], acc0); acc1 = fmaf(w0.y, f[1], acc1);
      acc0 = fmaf(w0.z, f[2], acc0); acc1 = fmaf(w0.w, f[3], acc1);
      acc0 = fmaf(w1.x, f[4], acc0); acc1 = fmaf(w1.y, f[5], acc1);
      acc0 = fmaf(w1.z, f[6], acc0); acc1 = fmaf(w1.w, f[7], acc1);
    }
  }
  const size_t o = ((size_t)n * a.H + y) * a.W + x;
  const float v = acc0 + acc1 + (a.residual ? __ldg(a.residual + o) : 0.f);
  a.y[o] = v;
  if (a.y8) {
    const P8Geom g1 = P8Geom::make(a.B, 1, a.H, a.W);
    float f[8] = {v, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    uint4 hi, lo;
    split8(f, hi, lo);
    *reinterpret_cast<uint4*>(a.y8 + g1.plane(n, 0, 0) + g1.pixel(y, x)) = hi;
    *reinterpret_cast<uint4*>(a.y8 + g1.plane(n, 0, 1) + g1.pixel(y, x)) = lo;
  }
}

template <int CI, int CO, int PXT>
static int launch_small(const SmallArgs& a, cudaStream_t st) {
  dim3 grid(ceil_div(a.W, SM_THREADS * PXT), a.H, a.B);
  char label[64];
  snprintf(label, sizeof(label), "small_to_p8 %d->%d @%dx%d", CI, CO, a.H, a.W);
  const double px = (double)a.B * a.H * a.W;
  {
    LaunchScope ls(st, label, 2.0 * CI * CO * (9 + (a.w1 ? 1 : 0)) * px, 4.0 * (CI + CO * (a.yq ? 2.0 : 1.0)) * px);
    small_to_p8_kernel<CI, CO, PXT><<<grid, SM_THREADS, 0, st>>>(a);
  }
  return (int)cudaGetLastError();
}

}  // namespace nconv

using namespace nconv;

extern "C" int nconv_b200_conv_small_p8(const float* x, const float* w, const float* w1, const float* scale,
                                        const float* shift, void* y, void* yq, int B, int CI, int CO, int H, int W,
                                        int relu, void* stream) {
  if (!x || !w || !y || B <= 0 || H <= 0 || W <= 0) return NCONV_ERR_BAD_ARG;
  if (yq && ((H | W) & 1)) return NCONV_ERR_UNSUPPORTED;
  SmallArgs a = {x, w, w1, scale, shift, (__nv_bfloat16*)y, (__nv_bfloat16*)yq, B, H, W, relu};
  cudaStream_t st = (cudaStream_t)stream;
  if (CI == 3 && CO == 32) return launch_small<3, 32, 2>(a, st);
  if (CI == 1 && CO == 32) return launch_small<1, 32, 2>(a, st);
  if (CI == 1 && CO == 64) return launch_small<1, 64, 1>(a, st);
  return NCONV_ERR_UNSUPPORTED;
}

extern "C" int nconv_b200_head_p8(const void* f, const float* w, const float* residual, float* y, void* y8, int B,
                                  int C, int H, int W, void* stream) {
  if (!f || !w || !y || B <= 0 || H <= 0 || W <= 0) return NCONV_ERR_BAD_ARG;
  if (C != 32 && C != 64) return NCONV_ERR_UNSUPPORTED;
  HeadArgs a = {(const __nv_bfloat16*)f, w, residual, y, (__nv_bfloat16*)y8, B, H, W};
  dim3 grid(ceil_div(W, 128), H, B);
  char label[64];
  snprintf(label, sizeof(label), "head_p8 %d->1 @%dx%d", C, H, W);
  const double px = (double)B * H * W;
  cudaStream_t st = (cudaStream_t)stream;
  LaunchScope ls(st, label, 2.0 * C * 9 * px, 4.0 * (C + 2 + (y8 ? 1 : 0)) * px);
  if (C == 64) head_p8_kernel<64><<<grid, 128, 0, st>>>(a);
  else head_p8_kernel<32><<<grid, 128, 0, st>>>(a);
  return (int)cudaGetLastError();
}
