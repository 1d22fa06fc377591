// Callers on either side of the training hot path:
//
//   * masked multi-resolution MSE:  /root/reference/utils.py:63-71 (calculate_loss_multi_resolution) on top
//     of :138-151 (calculate_loss, use_gradient_loss = False, the setting of train_step2.py:21,64):
//       up   = F.interpolate(d_s, size=(H, W), mode='bilinear', align_corners=False)
//       loss_s = mean over pixels of (up.masked_fill(gt == 0, 0) - gt)^2
//     The reference runs ~6 eager kernels per scale and materialises `up`; here one kernel per scale
//     interpolates on the fly and reduces, and the backward gathers the residual through the transposed
//     interpolation (deterministic: no atomics, fixed-order partial sums).
//   * AdamW over one flat parameter / gradient / moment buffer: utils.py:53-55 (optimizer.AdamW(lr, wd))
//     with torch.optim.AdamW's update rule (decoupled decay, bias-corrected moments).
#include "common.cuh"
#include "launch_count.h"
#include "nconv_b200.h"
#include <cstdint>
#include <cstdio>

namespace nconv {

// ATen's area_pixel_compute_source_index for align_corners=False (upsample_bilinear2d): negative sources clamp to 0
__device__ __forceinline__ void bilinear_src(int dst, float scale, int in_size, int& i0, int& i1, float& lam) {
  float src = ((float)dst + 0.5f) * scale - 0.5f;
  if (src < 0.f) src = 0.f;
  i0 = (int)src;
  if (i0 > in_size - 1) i0 = in_size - 1;
  i1 = i0 + ((i0 < in_size - 1) ? 1 : 0);
  lam = src - (float)i0;
}

__device__ __forceinline__ float upsample_at(const float* __restrict__ d, int h, int w, int y, int x, float sy, float sx) {
  int y0, y1, x0, x1; float ly, lx;
  bilinear_src(y, sy, h, y0, y1, ly);
  bilinear_src(x, sx, w, x0, x1, lx);
  const float a = __ldg(d + (size_t)y0 * w + x0), b = __ldg(d + (size_t)y0 * w + x1);
  const float c = __ldg(d + (size_t)y1 * w + x0), e = __ldg(d + (size_t)y1 * w + x1);
  // ATen: w0y*(w0x*a + w1x*b) + w1y*(w0x*c + w1x*e)
  return (1.f - ly) * ((1.f - lx) * a + lx * b) + ly * ((1.f - lx) * c + lx * e);
}

constexpr int LS_THREADS = 256;

static inline int grid_for(long long n, int threads) {
  long long b = (n + threads - 1) / threads;
  const long long cap = 148LL * 16;
  return (int)(b < 1 ? 1 : (b > cap ? cap : b));
}

// partial[block] = sum over the block's pixels of (mask * up - gt)^2   (double).  One thread = QV consecutive pixels of
// one row (QV = 4 with 16-byte loads when W % 4 == 0, else 1); IDENT: h == H, w == W (the interpolation is the identity).
template <int QV, bool IDENT>
__global__ void __launch_bounds__(LS_THREADS) masked_mse_fwd_kernel(const float* __restrict__ d, const float* __restrict__ gt,
                                                                    double* __restrict__ partial, int N, int h, int w, int H, int W) {
  __shared__ double s_red[LS_THREADS / 32];
  const float sy = (float)h / (float)H, sx = (float)w / (float)W;
  const int WQ = W / QV;
  const long long total = (long long)N * H * WQ;
  double acc = 0.0;
  for (long long q = (long long)blockIdx.x * LS_THREADS + threadIdx.x; q < total; q += (long long)gridDim.x * LS_THREADS) {
    const long long row = q / WQ;                       // n*H + y
    const int x = (int)(q - row * WQ) * QV;
    float g[QV], u[QV];
    if constexpr (QV == 4) {
      const float4 t = __ldg(reinterpret_cast<const float4*>(gt + row * W + x));
      g[0] = t.x; g[1] = t.y; g[2] = t.z; g[3] = t.w;
    } else {
      g[0] = __ldg(gt + row * W + x);
    }
    if constexpr (IDENT) {
      if constexpr (QV == 4) {
        const float4 t = __ldg(reinterpret_cast<const float4*>(d + row * W + x));
        u[0] = t.x; u[1] = t.y; u[2] = t.z; u[3] = t.w;
      } else {
        u[0] = __ldg(d + row * W + x);
      }
    } else {
      const int n = (int)(row / H), y = (int)(row - (long long)n * H);
      const float* dn = d + (size_t)n * h * w;
      int y0, y1; float ly;
      bilinear_src(y, sy, h, y0, y1, ly);
      const float* r0 = dn + (size_t)y0 * w; const float* r1 = dn + (size_t)y1 * w;
#pragma unroll
      for (int k = 0; k < QV; ++k) {
        int x0, x1; float lx;
        bilinear_src(x + k, sx, w, x0, x1, lx);
        u[k] = (1.f - ly) * ((1.f - lx) * __ldg(r0 + x0) + lx * __ldg(r0 + x1)) +
               ly * ((1.f - lx) * __ldg(r1 + x0) + lx * __ldg(r1 + x1));
      }
    }
    float part = 0.f;
#pragma unroll
    for (int k = 0; k < QV; ++k) {
      const float r = (g[k] != 0.f) ? u[k] - g[k] : 0.f;
      part = fmaf(r, r, part);
    }
    acc += (double)part;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int k = 0; k < LS_THREADS / 32; ++k) t += s_red[k];
    partial[blockIdx.x] = t;
  }
}

// loss[0] (+)= weight * sum(partial) / count      (one block, fixed order)
__global__ void masked_mse_final_kernel(const double* __restrict__ partial, int nblocks, double count, float weight,
                                        float* __restrict__ loss, int accumulate) {
  __shared__ double s_red[32];
  double t = 0.0;
  for (int i = threadIdx.x; i < nblocks; i += blockDim.x) t += partial[i];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
  if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = t;
  __syncthreads();
  if (threadIdx.x == 0) {
    double s = 0.0;
    for (int k = 0; k < (int)(blockDim.x >> 5); ++k) s += s_red[k];
    const float v = (float)(s / count) * weight;
    loss[0] = accumulate ? loss[0] + v : v;
  }
}

// R[n, y, x] = scale * 2 * mask * (up - gt): the loss gradient at the (virtual) upsampled prediction.  With IDENT this
// is the final gradient.  Same thread layout as the forward kernel.
template <int QV, bool IDENT>
__global__ void __launch_bounds__(LS_THREADS) masked_mse_resid_kernel(const float* __restrict__ d, const float* __restrict__ gt,
                                                                      const float* __restrict__ gloss, float* __restrict__ R,
                                                                      float coef, int N, int h, int w, int H, int W) {
  const float sy = (float)h / (float)H, sx = (float)w / (float)W;
  const int WQ = W / QV;
  const long long total = (long long)N * H * WQ;
  const float scale = 2.f * coef * __ldg(gloss);
  for (long long q = (long long)blockIdx.x * LS_THREADS + threadIdx.x; q < total; q += (long long)gridDim.x * LS_THREADS) {
    const long long row = q / WQ;
    const int x = (int)(q - row * WQ) * QV;
    float g[QV], u[QV];
    if constexpr (QV == 4) {
      const float4 t = __ldg(reinterpret_cast<const float4*>(gt + row * W + x));
      g[0] = t.x; g[1] = t.y; g[2] = t.z; g[3] = t.w;
    } else {
      g[0] = __ldg(gt + row * W + x);
    }
    if constexpr (IDENT) {
      if constexpr (QV == 4) {
        const float4 t = __ldg(reinterpret_cast<const float4*>(d + row * W + x));
        u[0] = t.x; u[1] = t.y; u[2] = t.z; u[3] = t.w;
      } else {
        u[0] = __ldg(d + row * W + x);
      }
    } else {
      const int n = (int)(row / H), y = (int)(row - (long long)n * H);
      const float* dn = d + (size_t)n * h * w;
      int y0, y1; float ly;
      bilinear_src(y, sy, h, y0, y1, ly);
      const float* r0 = dn + (size_t)y0 * w; const float* r1 = dn + (size_t)y1 * w;
#pragma unroll
      for (int k = 0; k < QV; ++k) {
        int x0, x1; float lx;
        bilinear_src(x + k, sx, w, x0, x1, lx);
        u[k] = (1.f - ly) * ((1.f - lx) * __ldg(r0 + x0) + lx * __ldg(r0 + x1)) +
               ly * ((1.f - lx) * __ldg(r1 + x0) + lx * __ldg(r1 + x1));
      }
    }
    float o[QV];
#pragma unroll
    for (int k = 0; k < QV; ++k) o[k] = (g[k] != 0.f) ? (u[k] - g[k]) * scale : 0.f;
    if constexpr (QV == 4) *reinterpret_cast<float4*>(R + row * W + x) = make_float4(o[0], o[1], o[2], o[3]);
    else R[row * W + x] = o[0];
  }
}

// ---- use_gradient_loss = True (utils.py:142-147, the setting of train_step1.py:20) -----------------------------
//   loss = 0.8 * sqrt(mse(rec, gt)) + 0.2 * (mean|sobel_x(gt - rec)| + mean|sobel_y(gt - rec)|),  rec = mask * up
// with zero-padded 3x3 cross-correlations (utils.py:95-136).  One CTA = one SB_TX x SB_TY tile of one sample.
constexpr int SB_TX = 32, SB_TY = 8, SB_THREADS = SB_TX * SB_TY;

// diff(y, x) = gt - mask*up inside the image, 0 outside (the convolution's zero padding)
template <bool IDENT>
__device__ __forceinline__ float sobel_diff(const float* __restrict__ dn, const float* __restrict__ gn, int h, int w, int H,
                                            int W, int y, int x, float sy, float sx) {
  if ((unsigned)y >= (unsigned)H || (unsigned)x >= (unsigned)W) return 0.f;
  const float g = __ldg(gn + (size_t)y * W + x);
  if (g == 0.f) return 0.f;
  if constexpr (IDENT) return g - __ldg(dn + (size_t)y * W + x);
  else return g - upsample_at(dn, h, w, y, x, sy, sx);
}

// partial[block] = { sum r^2, sum |gx|, sum |gy| } over the tile
template <bool IDENT>
__global__ void __launch_bounds__(SB_THREADS) sobel_loss_fwd_kernel(const float* __restrict__ d, const float* __restrict__ gt,
                                                                    double* __restrict__ partial, int h, int w, int H, int W) {
  __shared__ float s_d[SB_TY + 2][SB_TX + 2];
  __shared__ double s_red[3][SB_THREADS / 32];
  const int n = blockIdx.z, x0 = blockIdx.x * SB_TX, y0 = blockIdx.y * SB_TY;
  const float sy = (float)h / (float)H, sx = (float)w / (float)W;
  const float* dn = d + (size_t)n * h * w;
  const float* gn = gt + (size_t)n * H * W;
  for (int i = threadIdx.x; i < (SB_TY + 2) * (SB_TX + 2); i += SB_THREADS) {
    const int ty = i / (SB_TX + 2), tx = i % (SB_TX + 2);
    s_d[ty][tx] = sobel_diff<IDENT>(dn, gn, h, w, H, W, y0 + ty - 1, x0 + tx - 1, sy, sx);
  }
  __syncthreads();
  const int tx = threadIdx.x % SB_TX, ty = threadIdx.x / SB_TX;
  double v[3] = {0.0, 0.0, 0.0};
  if (y0 + ty < H && x0 + tx < W) {
    const float a00 = s_d[ty][tx], a01 = s_d[ty][tx + 1], a02 = s_d[ty][tx + 2];
    const float a10 = s_d[ty + 1][tx], a11 = s_d[ty + 1][tx + 1], a12 = s_d[ty + 1][tx + 2];
    const float a20 = s_d[ty + 2][tx], a21 = s_d[ty + 2][tx + 1], a22 = s_d[ty + 2][tx + 2];
    const float gx = (a00 - a02) + 2.f * (a10 - a12) + (a20 - a22);
    const float gy = (a00 - a20) + 2.f * (a01 - a21) + (a02 - a22);
    v[0] = (double)a11 * (double)a11; v[1] = fabsf(gx); v[2] = fabsf(gy);
  }
#pragma unroll
  for (int k = 0; k < 3; ++k) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v[k] += __shfl_xor_sync(0xffffffffu, v[k], o);
    if ((threadIdx.x & 31) == 0) s_red[k][threadIdx.x >> 5] = v[k];
  }
  __syncthreads();
  if (threadIdx.x < 3) {
    double t = 0.0;
    for (int k = 0; k < SB_THREADS / 32; ++k) t += s_red[threadIdx.x][k];
    partial[((size_t)(blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x) * 3 + threadIdx.x] = t;
  }
}

// stats[0..2] = the three sums; loss (+)= weight * (0.8*sqrt(S0/count) + 0.2*(S1+S2)/count)
__global__ void sobel_loss_final_kernel(const double* __restrict__ partial, int nblocks, double count, float weight,
                                        double* __restrict__ stats, float* __restrict__ loss, int accumulate) {
  __shared__ double s_red[3][32];
  double t[3] = {0.0, 0.0, 0.0};
  for (int i = threadIdx.x; i < nblocks; i += blockDim.x) {
    t[0] += partial[(size_t)i * 3]; t[1] += partial[(size_t)i * 3 + 1]; t[2] += partial[(size_t)i * 3 + 2];
  }
#pragma unroll
  for (int k = 0; k < 3; ++k) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) t[k] += __shfl_xor_sync(0xffffffffu, t[k], o);
    if ((threadIdx.x & 31) == 0) s_red[k][threadIdx.x >> 5] = t[k];
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    double s[3] = {0.0, 0.0, 0.0};
    for (int k = 0; k < (int)(blockDim.x >> 5); ++k) { s[0] += s_red[0][k]; s[1] += s_red[1][k]; s[2] += s_red[2][k]; }
    stats[0] = s[0]; stats[1] = s[1]; stats[2] = s[2];
    // the reference evaluates sqrt(mse) and the two means in fp32 and combines them in fp32
    const float v = ((float)sqrt(s[0] / count) * 0.8f + ((float)(s[1] / count) + (float)(s[2] / count)) * 0.2f) * weight;
    loss[0] = accumulate ? loss[0] + v : v;
  }
}

// R[n, y, x] = gloss * weight * dloss/d up(y, x): the gradient at the (virtual) upsampled prediction
template <bool IDENT>
__global__ void __launch_bounds__(SB_THREADS) sobel_loss_bwd_kernel(const float* __restrict__ d, const float* __restrict__ gt,
                                                                    const float* __restrict__ gloss,
                                                                    const double* __restrict__ stats, float* __restrict__ R,
                                                                    float weight, double count, int h, int w, int H, int W) {
  __shared__ float s_d[SB_TY + 4][SB_TX + 4];
  __shared__ float s_sx[SB_TY + 2][SB_TX + 2], s_sy[SB_TY + 2][SB_TX + 2];
  const int n = blockIdx.z, x0 = blockIdx.x * SB_TX, y0 = blockIdx.y * SB_TY;
  const float sy = (float)h / (float)H, sx = (float)w / (float)W;
  const float* dn = d + (size_t)n * h * w;
  const float* gn = gt + (size_t)n * H * W;
  for (int i = threadIdx.x; i < (SB_TY + 4) * (SB_TX + 4); i += SB_THREADS) {
    const int ty = i / (SB_TX + 4), tx = i % (SB_TX + 4);
    s_d[ty][tx] = sobel_diff<IDENT>(dn, gn, h, w, H, W, y0 + ty - 2, x0 + tx - 2, sy, sx);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < (SB_TY + 2) * (SB_TX + 2); i += SB_THREADS) {
    const int ty = i / (SB_TX + 2), tx = i % (SB_TX + 2);
    const int y = y0 + ty - 1, x = x0 + tx - 1;
    float vx = 0.f, vy = 0.f;
    if ((unsigned)y < (unsigned)H && (unsigned)x < (unsigned)W) {      // only in-image outputs enter the means
      const float a00 = s_d[ty][tx], a01 = s_d[ty][tx + 1], a02 = s_d[ty][tx + 2];
      const float a10 = s_d[ty + 1][tx], a12 = s_d[ty + 1][tx + 2];
      const float a20 = s_d[ty + 2][tx], a21 = s_d[ty + 2][tx + 1], a22 = s_d[ty + 2][tx + 2];
      const float gx = (a00 - a02) + 2.f * (a10 - a12) + (a20 - a22);
      const float gy = (a00 - a20) + 2.f * (a01 - a21) + (a02 - a22);
      vx = (gx > 0.f) ? 1.f : ((gx < 0.f) ? -1.f : 0.f);
      vy = (gy > 0.f) ? 1.f : ((gy < 0.f) ? -1.f : 0.f);
    }
    s_sx[ty][tx] = vx; s_sy[ty][tx] = vy;
  }
  __syncthreads();
  const int tx = threadIdx.x % SB_TX, ty = threadIdx.x / SB_TX;
  const int y = y0 + ty, x = x0 + tx;
  if (y >= H || x >= W) return;
  float out = 0.f;
  if (__ldg(gn + (size_t)y * W + x) != 0.f) {
    // d/d diff(q) of sum_p |gx(p)| = sum_{a,b} Kx[a][b] * sign(gx(q - (a-1, b-1))), same for y
    const float* X0 = &s_sx[ty][tx]; const float* Y0 = &s_sy[ty][tx];
    constexpr int P = SB_TX + 2;
    // p = q - (a-1, b-1)  ->  tile index (ty+1-(a-1), tx+1-(b-1)) = (ty+2-a, tx+2-b)
    const float gxs = (X0[2 * P + 2] - X0[2 * P]) + 2.f * (X0[P + 2] - X0[P]) + (X0[2] - X0[0]);
    const float gys = (Y0[2 * P + 2] - Y0[2]) + 2.f * (Y0[2 * P + 1] - Y0[1]) + (Y0[2 * P] - Y0[0]);
    const float diff = s_d[ty + 2][tx + 2];                                  // gt - up
    const float rmse = (float)sqrt(stats[0] / count);
    // d(0.8*sqrt(mse))/d up = 0.8 * (1/(2*rmse)) * 2*(up - gt)/count ;  d diff/d up = -1
    const float g_mse = (0.8f * 0.5f / rmse) * (2.f * (-diff) / (float)count);
    const float g_sob = -(0.2f / (float)count) * (gxs + gys);
    out = (g_mse + g_sob) * weight * __ldg(gloss);
  }
  R[((size_t)n * H + y) * W + x] = out;
}

// gd[n,i,j] = sum over high-resolution pixels of (transposed bilinear weight) * R   (R already masked and scaled).
// The contributing rows / columns are those whose source index lies in (i-1, i+1): a conservative window is narrowed
// to the exact column range once.  TPO lanes share one output: lane k owns columns xlo+k, xlo+k+TPO, ... (CPL of them,
// weights in registers), walks all rows, and the group is summed by a fixed-order butterfly (deterministic).
template <int TPO, int CPL>
__global__ void __launch_bounds__(LS_THREADS) bilinear_af_bwd_gather_kernel(const float* __restrict__ R, float* __restrict__ gd,
                                                                            int N, int h, int w, int H, int W) {
  const float sy = (float)h / (float)H, sx = (float)w / (float)W;
  const float fy = (float)H / (float)h, fx = (float)W / (float)w;
  const long long total = (long long)N * h * w;
  const int sub = threadIdx.x % TPO;
  const long long stride = (long long)gridDim.x * (LS_THREADS / TPO);
  // the loop bound is uniform over the CTA, so every lane reaches the butterfly below; lanes past the end compute
  // a clamped output and do not store it
  for (long long base = (long long)blockIdx.x * (LS_THREADS / TPO); base < total; base += stride) {
    const bool live = base + threadIdx.x / TPO < total;
    const long long idx = live ? base + threadIdx.x / TPO : total - 1;
    const int j = (int)(idx % w);
    const int i = (int)((idx / w) % h);
    const int n = (int)(idx / ((long long)w * h));
    const float* Rn = R + (size_t)n * H * W;
    int ylo = (int)floorf(((float)i - 0.5f) * fy - 0.5f) - 1, yhi = (int)ceilf(((float)i + 1.5f) * fy - 0.5f) + 1;
    int xlo = (int)floorf(((float)j - 0.5f) * fx - 0.5f) - 1, xhi = (int)ceilf(((float)j + 1.5f) * fx - 0.5f) + 1;
    ylo = max(ylo, 0); yhi = min(yhi, H - 1); xlo = max(xlo, 0); xhi = min(xhi, W - 1);
    auto wcol = [&](int x) { int x0, x1; float lx; bilinear_src(x, sx, w, x0, x1, lx);
                             return (x0 == j ? 1.f - lx : 0.f) + (x1 == j ? lx : 0.f); };
    auto touches = [&](int x) { int x0, x1; float lx; bilinear_src(x, sx, w, x0, x1, lx); return x0 == j || x1 == j; };
    while (xlo <= xhi && !touches(xlo)) ++xlo;
    while (xhi >= xlo && !touches(xhi)) --xhi;
    float wc[CPL]; int xc[CPL];
#pragma unroll
    for (int k = 0; k < CPL; ++k) {
      const int x = xlo + sub + k * TPO;
      wc[k] = (x <= xhi) ? wcol(x) : 0.f;
      xc[k] = min(x, W - 1);
    }
    float acc = 0.f;
    for (int y = ylo; y <= yhi; ++y) {
      int y0, y1; float ly;
      bilinear_src(y, sy, h, y0, y1, ly);
      const float wy = (y0 == i ? 1.f - ly : 0.f) + (y1 == i ? ly : 0.f);
      if (wy == 0.f) continue;
      const float* Rr = Rn + (size_t)y * W;
      float racc = 0.f;
#pragma unroll
      for (int k = 0; k < CPL; ++k) racc = fmaf(wc[k], __ldg(Rr + xc[k]), racc);
      for (int x = xlo + sub + CPL * TPO; x <= xhi; x += TPO) racc = fmaf(wcol(x), __ldg(Rr + x), racc);   // never in practice
      acc = fmaf(wy, racc, acc);
    }
#pragma unroll
    for (int o = TPO / 2; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (sub == 0 && live) gd[idx] = acc;
  }
}

static void launch_gather(cudaStream_t st, const float* R, float* gd, int N, int h, int w, int H, int W) {
  const long long lo = (long long)N * h * w;
  const int need = (int)(2.0 * (double)W / (double)w) + 2;          // columns with a source index inside (j-1, j+1)
  auto grid = [&](int tpo) { return grid_for(lo * tpo, LS_THREADS); };
  if (need <= 6) bilinear_af_bwd_gather_kernel<1, 6><<<grid(1), LS_THREADS, 0, st>>>(R, gd, N, h, w, H, W);
  else if (need <= 12) bilinear_af_bwd_gather_kernel<4, 3><<<grid(4), LS_THREADS, 0, st>>>(R, gd, N, h, w, H, W);
  else if (need <= 24) bilinear_af_bwd_gather_kernel<8, 3><<<grid(8), LS_THREADS, 0, st>>>(R, gd, N, h, w, H, W);
  else bilinear_af_bwd_gather_kernel<32, 3><<<grid(32), LS_THREADS, 0, st>>>(R, gd, N, h, w, H, W);
}

// torch.optim.AdamW (amsgrad=False, maximize=False), one element per thread over the flat buffers
__global__ void __launch_bounds__(256) adamw_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                                                    float* __restrict__ v, long long n, float lr, float beta1, float beta2,
                                                    float eps, float wd, float bc1, float bc2_sqrt) {
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n; i += (long long)gridDim.x * 256) {
    const float gi = g[i];
    float pi = p[i] * (1.f - lr * wd);
    const float mi = beta1 * m[i] + (1.f - beta1) * gi;          // exp_avg.lerp_(grad, 1 - beta1)
    const float vi = beta2 * v[i] + (1.f - beta2) * gi * gi;     // exp_avg_sq.mul_(beta2).addcmul_(grad, grad, 1 - beta2)
    m[i] = mi; v[i] = vi;
    const float denom = sqrtf(vi) / bc2_sqrt + eps;
    pi -= (lr / bc1) * (mi / denom);
    p[i] = pi;
  }
}

}  // namespace nconv

using namespace nconv;

extern "C" size_t nconv_b200_masked_mse_ws_bytes(int N, int h, int w, int H, int W) {
  if (N <= 0 || H <= 0 || W <= 0) return 0;
  const size_t partial = (size_t)148 * 16 * sizeof(double);
  const size_t resid = (h == H && w == W) ? 0 : (size_t)N * H * W * sizeof(float);
  return partial > resid ? partial : resid;
}

template <int QV, bool IDENT>
static void launch_mse_fwd(cudaStream_t st, int blocks, const float* d, const float* gt, double* ws, int N, int h, int w, int H,
                           int W) {
  masked_mse_fwd_kernel<QV, IDENT><<<blocks, LS_THREADS, 0, st>>>(d, gt, ws, N, h, w, H, W);
}
template <int QV, bool IDENT>
static void launch_mse_resid(cudaStream_t st, int blocks, const float* d, const float* gt, const float* gloss, float* R,
                             float coef, int N, int h, int w, int H, int W) {
  masked_mse_resid_kernel<QV, IDENT><<<blocks, LS_THREADS, 0, st>>>(d, gt, gloss, R, coef, N, h, w, H, W);
}
static inline bool quad_ok(const void* a, const void* b, const void* c, int W, bool ident) {
  return W % 4 == 0 && ((uintptr_t)a % 16) == 0 && (!ident || ((uintptr_t)b % 16) == 0) && (!c || ((uintptr_t)c % 16) == 0);
}

extern "C" int nconv_b200_masked_mse_bilinear_fwd(const float* d, const float* gt, float* loss, void* ws, size_t ws_bytes,
                                                  int N, int h, int w, int H, int W, float weight, int accumulate,
                                                  void* stream) {
  if (!d || !gt || !loss || !ws || N <= 0 || h <= 0 || w <= 0 || H <= 0 || W <= 0) return NCONV_ERR_BAD_ARG;
  if (ws_bytes < (size_t)148 * 16 * sizeof(double)) return NCONV_ERR_WORKSPACE;
  cudaStream_t st = (cudaStream_t)stream;
  const long long total = (long long)N * H * W;
  const bool ident = (h == H && w == W), quad = quad_ok(gt, d, nullptr, W, ident);
  const int blocks = grid_for(quad ? total / 4 : total, LS_THREADS);
  {
    LaunchScope ls(st, ident ? "masked_mse_fwd(ident)" : "masked_mse_fwd", 8.0 * total, 4.0 * (total + (double)N * h * w));
    if (quad) { if (ident) launch_mse_fwd<4, true>(st, blocks, d, gt, (double*)ws, N, h, w, H, W);
                else launch_mse_fwd<4, false>(st, blocks, d, gt, (double*)ws, N, h, w, H, W); }
    else { if (ident) launch_mse_fwd<1, true>(st, blocks, d, gt, (double*)ws, N, h, w, H, W);
           else launch_mse_fwd<1, false>(st, blocks, d, gt, (double*)ws, N, h, w, H, W); }
  }
  NCONV_RETURN_IF_ERR(cudaGetLastError());
  {
    LaunchScope ls(st, "masked_mse_final", 0.0, 8.0 * blocks);
    masked_mse_final_kernel<<<1, 256, 0, st>>>((const double*)ws, blocks, (double)total, weight, loss, accumulate);
  }
  return (int)cudaGetLastError();
}

extern "C" int nconv_b200_masked_mse_bilinear_bwd(const float* d, const float* gt, const float* gloss, float* gd, void* ws,
                                                  size_t ws_bytes, int N, int h, int w, int H, int W, float weight,
                                                  void* stream) {
  if (!d || !gt || !gloss || !gd || N <= 0 || h <= 0 || w <= 0 || H <= 0 || W <= 0) return NCONV_ERR_BAD_ARG;
  const bool ident = (h == H && w == W);
  if (!ident && (!ws || ws_bytes < (size_t)N * H * W * sizeof(float))) return NCONV_ERR_WORKSPACE;
  cudaStream_t st = (cudaStream_t)stream;
  const long long total = (long long)N * H * W, lo = (long long)N * h * w;
  const float coef = weight / (float)((double)N * H * W);
  float* R = ident ? gd : (float*)ws;
  const bool quad = quad_ok(gt, d, R, W, ident);
  const int blocks = grid_for(quad ? total / 4 : total, LS_THREADS);
  {
    LaunchScope ls(st, ident ? "masked_mse_bwd(ident)" : "masked_mse_resid", 8.0 * total, 4.0 * (2.0 * total + (double)lo));
    if (quad) { if (ident) launch_mse_resid<4, true>(st, blocks, d, gt, gloss, R, coef, N, h, w, H, W);
                else launch_mse_resid<4, false>(st, blocks, d, gt, gloss, R, coef, N, h, w, H, W); }
    else { if (ident) launch_mse_resid<1, true>(st, blocks, d, gt, gloss, R, coef, N, h, w, H, W);
           else launch_mse_resid<1, false>(st, blocks, d, gt, gloss, R, coef, N, h, w, H, W); }
  }
  NCONV_RETURN_IF_ERR(cudaGetLastError());
  if (!ident) {
    LaunchScope ls(st, "bilinear_af_bwd_gather", 8.0 * total, 4.0 * (total + lo));
    launch_gather(st, R, gd, N, h, w, H, W);
  }
  return (int)cudaGetLastError();
}

extern "C" int nconv_b200_adamw_step(float* p, const float* g, float* m, float* v, long long n, float lr, float beta1,
                                     float beta2, float eps, float weight_decay, int step, void* stream) {
  if (!p || !g || !m || !v || n <= 0 || step < 1) return NCONV_ERR_BAD_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  const double bc1 = 1.0 - pow((double)beta1, (double)step), bc2 = 1.0 - pow((double)beta2, (double)step);
  LaunchScope ls(st, "adamw_step", 12.0 * n, 28.0 * n);
  adamw_kernel<<<grid_for(n, 256), 256, 0, st>>>(p, g, m, v, n, lr, beta1, beta2, eps, weight_decay, (float)bc1,
                                                 (float)sqrt(bc2));
  return (int)cudaGetLastError();
}

// ---- use_gradient_loss = True ------------------------------------------------------------------------------
static inline dim3 sobel_grid(int N, int H, int W) { return dim3((W + SB_TX - 1) / SB_TX, (H + SB_TY - 1) / SB_TY, N); }

extern "C" size_t nconv_b200_sobel_loss_ws_bytes(int N, int h, int w, int H, int W) {
  if (N <= 0 || H <= 0 || W <= 0) return 0;
  const dim3 g = sobel_grid(N, H, W);
  const size_t partial = (size_t)g.x * g.y * g.z * 3 * sizeof(double);
  const size_t resid = (h == H && w == W) ? 0 : (size_t)N * H * W * sizeof(float);
  return partial > resid ? partial : resid;
}

extern "C" int nconv_b200_sobel_loss_fwd(const float* d, const float* gt, float* loss, double* stats, void* ws,
                                         size_t ws_bytes, int N, int h, int w, int H, int W, float weight, int accumulate,
                                         void* stream) {
  if (!d || !gt || !loss || !stats || !ws || N <= 0 || h <= 0 || w <= 0 || H <= 0 || W <= 0) return NCONV_ERR_BAD_ARG;
  if (N > 65535) return NCONV_ERR_UNSUPPORTED;
  if (ws_bytes < nconv_b200_sobel_loss_ws_bytes(N, h, w, H, W)) return NCONV_ERR_WORKSPACE;
  cudaStream_t st = (cudaStream_t)stream;
  const dim3 g = sobel_grid(N, H, W);
  const double total = (double)N * H * W;
  {
    LaunchScope ls(st, "sobel_loss_fwd", 30.0 * total, 4.0 * (total + (double)N * h * w));
    if (h == H && w == W) sobel_loss_fwd_kernel<true><<<g, SB_THREADS, 0, st>>>(d, gt, (double*)ws, h, w, H, W);
    else sobel_loss_fwd_kernel<false><<<g, SB_THREADS, 0, st>>>(d, gt, (double*)ws, h, w, H, W);
  }
  NCONV_RETURN_IF_ERR(cudaGetLastError());
  {
    LaunchScope ls(st, "sobel_loss_final", 0.0, 24.0 * g.x * g.y * g.z);
    sobel_loss_final_kernel<<<1, 1024, 0, st>>>((const double*)ws, (int)(g.x * g.y * g.z), total, weight, stats, loss, accumulate);
  }
  return (int)cudaGetLastError();
}

extern "C" int nconv_b200_sobel_loss_bwd(const float* d, const float* gt, const float* gloss, const double* stats, float* gd,
                                         void* ws, size_t ws_bytes, int N, int h, int w, int H, int W, float weight,
                                         void* stream) {
  if (!d || !gt || !gloss || !stats || !gd || N <= 0 || h <= 0 || w <= 0 || H <= 0 || W <= 0) return NCONV_ERR_BAD_ARG;
  if (N > 65535) return NCONV_ERR_UNSUPPORTED;
  const bool same = (h == H && w == W);
  if (!same && (!ws || ws_bytes < (size_t)N * H * W * sizeof(float))) return NCONV_ERR_WORKSPACE;
  cudaStream_t st = (cudaStream_t)stream;
  const double total = (double)N * H * W;
  float* R = same ? gd : (float*)ws;
  {
    LaunchScope ls(st, "sobel_loss_bwd", 60.0 * total, 4.0 * (2.0 * total + (double)N * h * w));
    if (same) sobel_loss_bwd_kernel<true><<<sobel_grid(N, H, W), SB_THREADS, 0, st>>>(d, gt, gloss, stats, R, weight, total, h, w, H, W);
    else sobel_loss_bwd_kernel<false><<<sobel_grid(N, H, W), SB_THREADS, 0, st>>>(d, gt, gloss, stats, R, weight, total, h, w, H, W);
  }
  NCONV_RETURN_IF_ERR(cudaGetLastError());
  if (!same) {
    const long long lo = (long long)N * h * w;
    LaunchScope ls(st, "bilinear_af_bwd_gather", 8.0 * total, 4.0 * (total + lo));
    launch_gather(st, R, gd, N, h, w, H, W);
  }
  return (int)cudaGetLastError();
}
