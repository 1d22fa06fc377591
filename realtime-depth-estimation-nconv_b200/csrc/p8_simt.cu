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
constexpr int SM_ROWS = 8;        // rows per CTA of small_to_p8

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
  // Input tile of the CTA's SM_ROWS rows (+1 halo each side), zero-filled outside the image, copied global -> shared
  // once with 4-byte cp.async (rows start at x0-1: no wider alignment).  Before this the 27 loads per pixel were
  // issued row by row and ncu showed 28 % of the warp samples stalled on them (long scoreboard at the first FFMA2 of
  // every input channel); the 158 registers of the accumulators leave no room for a register prefetch.
  // (The one-pixel-per-thread variants, 1->64, measured slower with the tile: they keep the direct loads.)
  constexpr bool STAGE = PXT > 1;
  constexpr int TWX = SM_THREADS * PXT + 2;
  __shared__ float s_in[STAGE ? CI : 1][STAGE ? SM_ROWS + 2 : 1][STAGE ? TWX : 1];
  const int n = blockIdx.z;
  if constexpr (STAGE) {
    const int x0 = blockIdx.x * (SM_THREADS * PXT) - 1, y0 = blockIdx.y * SM_ROWS - 1;
#pragma unroll 4
    for (int i = tid; i < CI * (SM_ROWS + 2) * TWX; i += SM_THREADS) {
      const int c = i % TWX, r = (i / TWX) % (SM_ROWS + 2), ci = i / (TWX * (SM_ROWS + 2));
      const int yy = y0 + r, xx = x0 + c;
      const bool ok = yy >= 0 && yy < a.H && xx >= 0 && xx < a.W;
      const float* src = a.x + (ok ? (((size_t)n * CI + ci) * a.H + yy) * a.W + xx : 0);
      const unsigned dst = (unsigned)__cvta_generic_to_shared(&s_in[ci][r][c]);
      const int nbytes = ok ? 4 : 0;
      asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(dst), "l"(src), "r"(nbytes) : "memory");
    }
    asm volatile("cp.async.wait_all;" ::: "memory");
  }
  __syncthreads();

  const int xb = blockIdx.x * (SM_THREADS * PXT) + tid;
  const P8Geom g = P8Geom::make(a.B, CO, a.H, a.W);
  const P8QGeom qg = P8QGeom::make(a.B, CO, a.H, a.W);
  // a CTA walks SM_ROWS rows of its column block: the weight staging above is paid once per SM_ROWS rows
#pragma unroll 1
  for (int y = blockIdx.y * SM_ROWS; y < min((int)(blockIdx.y + 1) * SM_ROWS, a.H); ++y) {
  // the weights are loop-invariant shared-memory loads: an opaque zero keeps the compiler from hoisting all
  // CI*9*CO of them out of the row loop into registers (255 registers + 3 KB of spills otherwise)
  int wz = 0;
  asm volatile("" : "+r"(wz));
  u64 acc2[PXT][CO / 2];          // packed pairs of output channels (FFMA2: weight pair x broadcast input)
#pragma unroll
  for (int p = 0; p < PXT; ++p)
#pragma unroll
    for (int o = 0; o < CO / 2; ++o) acc2[p][o] = 0ull;
  float centre[PXT][CI];

#pragma unroll
  for (int ci = 0; ci < CI; ++ci) {
    const float* plane = a.x + ((size_t)n * CI + ci) * a.H * a.W;
#pragma unroll
    for (int u = 0; u < 3; ++u) {
      const int r = y - blockIdx.y * SM_ROWS + u;          // tile row of image row y + u - 1
      const int yy = y + u - 1;
      float in[PXT][3];
#pragma unroll
      for (int p = 0; p < PXT; ++p)
#pragma unroll
        for (int v = 0; v < 3; ++v) {
          if constexpr (STAGE) {
            in[p][v] = s_in[ci][r][tid + p * SM_THREADS + v];
          } else {
            const int xx = xb + p * SM_THREADS + v - 1;
            in[p][v] = (yy >= 0 && yy < a.H && xx >= 0 && xx < a.W) ? __ldg(plane + (size_t)yy * a.W + xx) : 0.f;
          }
        }
      if (u == 1) {
#pragma unroll
        for (int p = 0; p < PXT; ++p) centre[p][ci] = in[p][1];
      }
#pragma unroll
      for (int v = 0; v < 3; ++v) {
        const float* wr = s_w + wz + (ci * 9 + u * 3 + v) * CO;
#pragma unroll
        for (int o = 0; o < CO; o += 4) {
          const ulonglong2 w4 = *reinterpret_cast<const ulonglong2*>(wr + o);
#pragma unroll
          for (int p = 0; p < PXT; ++p) {
            ffma2(acc2[p][o / 2], w4.x, in[p][v]);
            ffma2(acc2[p][o / 2 + 1], w4.y, in[p][v]);
          }
        }
      }
    }
  }

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
        float v = fmaf((o & 1) ? hi32f(acc2[p][o >> 1]) : lo32f(acc2[p][o >> 1]), s_scale[o], s_shift[o]);
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
}

struct HeadArgs {
  const __nv_bfloat16* f; const float* w; const float* residual; float* y; __nv_bfloat16* y8;
  int B, H, W;
};

// 3x3 C->1 head as "1x1 to nine planes, then a shifted sum":  t[tap][p] = sum_c w[c][tap] * f[c][p] is
// computed ONCE per pixel of the tile + halo (each feature vector is loaded and converted from split bf16
// once: 8 x 16-byte loads, instead of nine times through L1 as a direct 3x3 gather would), kept in shared
// memory, and out[y][x] = sum_{u,v} t[3u+v][y+u-1][x+v-1] + residual.  The P8 ring supplies the zero padding.
// FMAs are packed over tap pairs (FFMA2: weight pair from shared memory, the feature as the broadcast scalar).
constexpr int HD_TW = 64, HD_TH = 16, HD_PW = HD_TW + 2, HD_PH = HD_TH + 2, HD_THREADS = 256;

template <int C>
__global__ void __launch_bounds__(HD_THREADS) head_p8_kernel(const HeadArgs a) {
  __shared__ __align__(16) float s_w[C * 10];            // [c][tap 0..8, pad]
  __shared__ float s_t[9][HD_PH][HD_PW + 1];
  for (int i = threadIdx.x; i < C * 10; i += HD_THREADS) {
    const int tap = i % 10, c = i / 10;
    s_w[i] = tap < 9 ? __ldg(a.w + (size_t)c * 9 + tap) : 0.f;
  }
  __syncthreads();
  const int n = blockIdx.z, y0 = blockIdx.y * HD_TH, x0 = blockIdx.x * HD_TW;
  const P8Geom g = P8Geom::make(a.B, C, a.H, a.W);
  // ---- phase 1: nine partial dot products per staged pixel (padded coords: tile pixel (r, c) = image (y0+r-1, x0+c-1))
  // Two staged pixels per thread and pass: every weight pair read from shared memory (5 LDS.64 per channel) feeds two
  // FFMA2 instead of one — the kernel was L1/shared-pipe bound (ncu: l1tex 82 %) on exactly those broadcast reads.
  constexpr int HD_NP = 2;
  for (int p0 = threadIdx.x; p0 < HD_PH * HD_PW; p0 += HD_NP * HD_THREADS) {
    int r[HD_NP], c[HD_NP];
    bool live[HD_NP], in[HD_NP];
    size_t off[HD_NP];
    u64 acc[HD_NP][5];
#pragma unroll
    for (int k = 0; k < HD_NP; ++k) {
      const int p = p0 + k * HD_THREADS;
      live[k] = p < HD_PH * HD_PW;
      r[k] = p / HD_PW; c[k] = p - r[k] * HD_PW;
      const int py = y0 + r[k], px = x0 + c[k];          // padded plane coordinates
      in[k] = live[k] && py < a.H + 2 && px < a.W + 2;
      off[k] = in[k] ? ((size_t)py * g.Wp + px) * 8 : 0;
#pragma unroll
      for (int q = 0; q < 5; ++q) acc[k][q] = 0ull;
    }
#pragma unroll 1
    for (int ch = 0; ch < C / 8; ++ch) {
      float f[HD_NP][8];
#pragma unroll
      for (int k = 0; k < HD_NP; ++k) {
        const uint4 hi = __ldg(reinterpret_cast<const uint4*>(a.f + g.plane(n, ch, 0) + off[k]));
        const uint4 lo = __ldg(reinterpret_cast<const uint4*>(a.f + g.plane(n, ch, 1) + off[k]));
        join8(hi, lo, f[k]);
      }
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const float* wr = s_w + (ch * 8 + e) * 10;
#pragma unroll
        for (int q = 0; q < 5; ++q) {
          const u64 w2 = *reinterpret_cast<const u64*>(wr + 2 * q);
#pragma unroll
          for (int k = 0; k < HD_NP; ++k) ffma2(acc[k][q], w2, f[k][e]);
        }
      }
    }
#pragma unroll
    for (int k = 0; k < HD_NP; ++k) {
      if (!live[k]) continue;
#pragma unroll
      for (int q = 0; q < 5; ++q) {
        s_t[2 * q][r[k]][c[k]] = in[k] ? lo32f(acc[k][q]) : 0.f;
        if (2 * q + 1 < 9) s_t[2 * q + 1][r[k]][c[k]] = in[k] ? hi32f(acc[k][q]) : 0.f;
      }
    }
  }
  __syncthreads();
  // ---- phase 2: shifted sum + residual; 4 outputs per thread
  const P8Geom g1 = P8Geom::make(a.B, 1, a.H, a.W);
  for (int p = threadIdx.x; p < HD_TH * HD_TW; p += HD_THREADS) {
    const int r = p / HD_TW, c = p - r * HD_TW;
    const int y = y0 + r, x = x0 + c;
    if (y >= a.H || x >= a.W) continue;
    float v = 0.f;
#pragma unroll
    for (int u = 0; u < 3; ++u)
#pragma unroll
      for (int w = 0; w < 3; ++w) v += s_t[u * 3 + w][r + u][c + w];
    const size_t o = ((size_t)n * a.H + y) * a.W + x;
    if (a.residual) v += __ldg(a.residual + o);
    a.y[o] = v;
    if (a.y8) {
      float f[8] = {v, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      uint4 hi, lo;
      split8(f, hi, lo);
      *reinterpret_cast<uint4*>(a.y8 + g1.plane(n, 0, 0) + g1.pixel(y, x)) = hi;
      *reinterpret_cast<uint4*>(a.y8 + g1.plane(n, 0, 1) + g1.pixel(y, x)) = lo;
    }
  }
}

template <int CI, int CO, int PXT>
static int launch_small(const SmallArgs& a, cudaStream_t st) {
  dim3 grid(ceil_div(a.W, SM_THREADS * PXT), ceil_div(a.H, SM_ROWS), a.B);
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
  dim3 grid(ceil_div(W, HD_TW), ceil_div(H, HD_TH), B);
  char label[64];
  snprintf(label, sizeof(label), "head_p8 %d->1 @%dx%d", C, H, W);
  const double px = (double)B * H * W;
  cudaStream_t st = (cudaStream_t)stream;
  LaunchScope ls(st, label, 2.0 * C * 9 * px, 4.0 * (C + 2 + (y8 ? 1 : 0)) * px);
  if (C == 64) head_p8_kernel<64><<<grid, HD_THREADS, 0, st>>>(a);
  else head_p8_kernel<32><<<grid, HD_THREADS, 0, st>>>(a);
  return (int)cudaGetLastError();
}
