// P8 layout utilities (csrc/p8.cuh): fp32 NCHW <-> P8 conversion at the boundaries of the
// RGB-guided stack and in tests, and the zero ring of reused buffers.
#include "common.cuh"
#include "launch_count.h"
#include "nconv_b200.h"
#include "p8.cuh"

namespace nconv {

// fp32 NCHW -> P8.  One thread = one ring-padded pixel x the 8 channels of a chunk; grid (columns / 128, row groups, B * C8)
// so no index needs a division.  Optional fusions for the training backward (ops.py::_ConvTrain.backward):
//   mask  g = x * [mask > 0]      (ReLU backward: x is the output gradient, mask the saved activation)
//   acc   acc[c] += sum over pixels of the packed value (bias gradient), fp32 partials per CTA, fp64 atomics.
constexpr int PK_THREADS = 128, PK_ROWS = 8;

template <bool MASK>
__global__ void __launch_bounds__(PK_THREADS) p8_pack_kernel(const float* __restrict__ x, const float* __restrict__ mask,
                                                             __nv_bfloat16* __restrict__ p8, double* __restrict__ acc,
                                                             int B, int C, int H, int W) {
  const P8Geom g = P8Geom::make(B, C, H, W);
  const int xx = blockIdx.x * PK_THREADS + threadIdx.x;          // padded column
  const int ch = blockIdx.z % g.C8, n = blockIdx.z / g.C8;
  const int y0 = blockIdx.y * PK_ROWS;
  float sum[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) sum[e] = 0.f;
  if (xx < g.Wp) {
#pragma unroll 2
    for (int yy = y0; yy < y0 + PK_ROWS && yy < g.Hp; ++yy) {
      float f[8], m[8];
      const bool inside = yy >= 1 && yy <= H && xx >= 1 && xx <= W;
      // all loads of the pixel first (independent, issued back to back), then the mask
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const int c = ch * 8 + e;
        const bool ok = inside && c < C;
        const size_t i = ok ? (((size_t)n * C + c) * H + (yy - 1)) * W + (xx - 1) : 0;
        f[e] = ok ? __ldg(x + i) : 0.f;
        m[e] = (MASK && ok) ? __ldg(mask + i) : 1.f;
      }
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        if (MASK && !(m[e] > 0.f)) f[e] = 0.f;
        sum[e] += f[e];
      }
      uint4 hi, lo;
      split8_fast(f, hi, lo);
      const size_t off = ((size_t)yy * g.Wp + xx) * 8;
      *reinterpret_cast<uint4*>(p8 + g.plane(n, ch, 0) + off) = hi;
      *reinterpret_cast<uint4*>(p8 + g.plane(n, ch, 1) + off) = lo;
    }
  }
  if (acc) {
    __shared__ float sh[PK_THREADS / 32][8];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      float v = sum[e];
#pragma unroll
      for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
      if (lane == 0) sh[warp][e] = v;
    }
    __syncthreads();
    if (threadIdx.x < 8 && ch * 8 + threadIdx.x < C) {
      double t = 0.0;
#pragma unroll
      for (int w = 0; w < PK_THREADS / 32; ++w) t += (double)sh[w][threadIdx.x];
      atomicAdd(acc + ch * 8 + threadIdx.x, t);
    }
  }
}

// fp32 NCHW -> P8 from a channel sub-range and / or with the 2x2 spatial phases moved to channels (training backward of the
// transposed convolutions, ops.py::_ConvTTrain.backward):
//   PHASES = false: p8[n][c][i][j]              = x[n][c0 + c][i][j]                       x: [B][Ctot][H][W]
//   PHASES = true : p8[n][(py*2+px)*C + c][i][j] = x[n][c0 + c][2i + py][2j + px]           x: [B][Ctot][2H][2W], p8: 4C channels
// (C a multiple of 8 with PHASES, so a chunk of 8 output channels lies in one phase.)
template <bool PHASES>
__global__ void __launch_bounds__(PK_THREADS) p8_pack_ex_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ p8, int B, int C,
                                                                int H, int W, int Ctot, int c0) {
  const int CP = PHASES ? 4 * C : C;                             // channels of the P8 tensor
  const P8Geom g = P8Geom::make(B, CP, H, W);
  const int xx = blockIdx.x * PK_THREADS + threadIdx.x;          // padded column
  const int ch = blockIdx.z % g.C8, n = blockIdx.z / g.C8;
  const int y0 = blockIdx.y * PK_ROWS;
  if (xx >= g.Wp) return;
  const int ph = PHASES ? (ch * 8) / C : 0, cb = PHASES ? ch * 8 - ph * C : ch * 8;
  const int py = ph >> 1, px = ph & 1;
  const int Hs = PHASES ? 2 * H : H, Ws = PHASES ? 2 * W : W;
#pragma unroll 2
  for (int yy = y0; yy < y0 + PK_ROWS && yy < g.Hp; ++yy) {
    float f[8];
    const bool inside = yy >= 1 && yy <= H && xx >= 1 && xx <= W;
    const int sy = PHASES ? 2 * (yy - 1) + py : yy - 1, sx = PHASES ? 2 * (xx - 1) + px : xx - 1;
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int c = cb + e;
      const bool ok = inside && c < C;
      f[e] = ok ? __ldg(x + (((size_t)n * Ctot + c0 + c) * Hs + sy) * Ws + sx) : 0.f;
    }
    uint4 hi, lo;
    split8_fast(f, hi, lo);
    const size_t off = ((size_t)yy * g.Wp + xx) * 8;
    *reinterpret_cast<uint4*>(p8 + g.plane(n, ch, 0) + off) = hi;
    *reinterpret_cast<uint4*>(p8 + g.plane(n, ch, 1) + off) = lo;
  }
}

// ConvTranspose2d(4, 2, 1) as a 3x3 / stride-1 convolution over the phase-split output gradient (see ops.py):
//   w3[ci][(py*2+px)*CO + co][r][s] = w[ci][co][u][v]  with  py = 1: r = 0 -> u = 0, r = 1 -> u = 2;  py = 0: r = 1 -> u = 1,
//   r = 2 -> u = 3 (same for px, s, v); every other entry 0.   w: [CI][CO][4][4], w3: [CI][4*CO][3][3]
__device__ __forceinline__ int ctp_tap(int p, int r) { return p ? (r == 0 ? 0 : (r == 1 ? 2 : -1)) : (r == 1 ? 1 : (r == 2 ? 3 : -1)); }

__global__ void convT_phase_weights_kernel(const float* __restrict__ w, float* __restrict__ w3, int CI, int CO) {
  const int total = CI * 4 * CO * 9;
  for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += gridDim.x * blockDim.x) {
    const int s = idx % 3, r = (idx / 3) % 3;
    int q = idx / 9;
    const int co = q % CO; q /= CO;
    const int ph = q % 4, ci = q / 4;
    const int u = ctp_tap(ph >> 1, r), v = ctp_tap(ph & 1, s);
    w3[idx] = (u >= 0 && v >= 0) ? __ldg(w + (((size_t)ci * CO + co) * 4 + u) * 4 + v) : 0.f;
  }
}

// the inverse gather for the weight gradient: gw[lead + ci][co][u][v] = gw3[ci][(py*2+px)*CO + co][r][s]
// gw3: [CIF][4*CO][3][3] (CIF = the CI - lead feature channels), gw: [CIF + lead][CO][4][4]
__global__ void convT_phase_wgrad_kernel(const float* __restrict__ gw3, float* __restrict__ gw, int CIF, int CO, int lead) {
  const int total = CIF * CO * 16;
  for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += gridDim.x * blockDim.x) {
    const int v = idx & 3, u = (idx >> 2) & 3;
    const int co = (idx >> 4) % CO, ci = (idx >> 4) / CO;
    const int py = (u & 1) ? 0 : 1, px = (v & 1) ? 0 : 1;            // u = 0, 2: odd rows (py = 1); u = 1, 3: even rows
    const int r = py ? (u == 0 ? 0 : 1) : (u == 1 ? 1 : 2), s = px ? (v == 0 ? 0 : 1) : (v == 1 ? 1 : 2);
    gw[(((size_t)(lead + ci)) * CO + co) * 16 + u * 4 + v] = __ldg(gw3 + (((size_t)ci * 4 * CO + (py * 2 + px) * CO + co) * 3 + r) * 3 + s);
  }
}

// 3x3 / stride-2 / pad-1 Conv2d backward through the same phase split (include/nconv_b200.h):
//   w3d[(a*2+b)*CI + ci][co][r][s] = w[co][ci][u][v]   a = 0: r = 1 -> u = 1;  a = 1: r = 1 -> u = 2, r = 2 -> u = 0
__device__ __forceinline__ int cs2_dtap(int a, int r) { return a ? (r == 1 ? 2 : (r == 2 ? 0 : -1)) : (r == 1 ? 1 : -1); }

__global__ void convS2_phase_weights_kernel(const float* __restrict__ w, float* __restrict__ w3d, int CO, int CI) {
  const int total = 4 * CI * CO * 9;
  for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += gridDim.x * blockDim.x) {
    const int s = idx % 3, r = (idx / 3) % 3;
    int q = idx / 9;
    const int co = q % CO; q /= CO;
    const int ci = q % CI, ph = q / CI;
    const int u = cs2_dtap(ph >> 1, r), v = cs2_dtap(ph & 1, s);
    w3d[idx] = (u >= 0 && v >= 0) ? __ldg(w + (((size_t)co * CI + ci) * 3 + u) * 3 + v) : 0.f;
  }
}

//   gw[co][ci][u][v] = gw3[co][(py*2+px)*CI + ci][r][s]   u = 0: py = 1, r = 0;  u = 1: py = 0, r = 1;  u = 2: py = 1, r = 1
__global__ void convS2_phase_wgrad_kernel(const float* __restrict__ gw3, float* __restrict__ gw, int CO, int CI) {
  const int total = CO * CI * 9;
  for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += gridDim.x * blockDim.x) {
    const int v = idx % 3, u = (idx / 3) % 3;
    const int ci = (idx / 9) % CI, co = idx / (9 * CI);
    const int py = (u == 1) ? 0 : 1, px = (v == 1) ? 0 : 1;
    const int r = (u == 0) ? 0 : 1, s = (v == 0) ? 0 : 1;
    gw[idx] = __ldg(gw3 + (((size_t)co * 4 * CI + (py * 2 + px) * CI + ci) * 3 + r) * 3 + s);
  }
}

// P8 -> fp32 NCHW; the C channels land at channel offset c0 of a [B, Ctot, H, W] tensor (Ctot = C, c0 = 0: a plain unpack)
__global__ void __launch_bounds__(PK_THREADS) p8_unpack_kernel(const __nv_bfloat16* __restrict__ p8, float* __restrict__ x, int B, int C,
                                                               int H, int W, int Ctot, int c0) {
  const P8Geom g = P8Geom::make(B, C, H, W);
  const int xx = blockIdx.x * PK_THREADS + threadIdx.x;          // interior column
  const int ch = blockIdx.z % g.C8, n = blockIdx.z / g.C8;
  if (xx >= W) return;
  const int y0 = blockIdx.y * PK_ROWS;
  for (int yy = y0; yy < y0 + PK_ROWS && yy < H; ++yy) {
    const size_t off = g.pixel(yy, xx);
    const uint4 hi = *reinterpret_cast<const uint4*>(p8 + g.plane(n, ch, 0) + off);
    const uint4 lo = *reinterpret_cast<const uint4*>(p8 + g.plane(n, ch, 1) + off);
    float f[8];
    join8(hi, lo, f);
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int c = ch * 8 + e;
      if (c < C) x[(((size_t)n * Ctot + c0 + c) * H + yy) * W + xx] = f[e];
    }
  }
}

// zero ring (rows 0 and Hp-1, columns 0 and Wp-1) of every plane
__global__ void p8_zero_ring_kernel(__nv_bfloat16* __restrict__ p8, int planes, int Hp, int Wp) {
  const int per = 2 * Wp + 2 * (Hp - 2);
  const size_t total = (size_t)planes * per;
  const uint4 z = make_uint4(0, 0, 0, 0);
  for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
    int i = (int)(idx % per);
    size_t pl = idx / per;
    int y, x;
    if (i < Wp) { y = 0; x = i; }
    else if (i < 2 * Wp) { y = Hp - 1; x = i - Wp; }
    else { int k = i - 2 * Wp; y = 1 + (k >> 1); x = (k & 1) ? Wp - 1 : 0; }
    *reinterpret_cast<uint4*>(p8 + (pl * Hp * Wp + (size_t)y * Wp + x) * 8) = z;
  }
}

}  // namespace nconv

using namespace nconv;

extern "C" size_t nconv_b200_p8_bytes(int B, int C, int H, int W) {
  if (B <= 0 || C <= 0 || H <= 0 || W <= 0) return 0;
  return P8Geom::make(B, C, H, W).total_elems() * sizeof(__nv_bfloat16) + P8_TAIL_SLACK_BYTES;
}

static int p8_pack_launch(const float* x, const float* mask, void* p8, double* acc, int B, int C, int H, int W, const char* label,
                          cudaStream_t st) {
  const P8Geom g = P8Geom::make(B, C, H, W);
  if ((long long)B * g.C8 > 65535) return NCONV_ERR_UNSUPPORTED;
  dim3 grid((g.Wp + PK_THREADS - 1) / PK_THREADS, (g.Hp + PK_ROWS - 1) / PK_ROWS, B * g.C8);
  const double n = (double)B * C * H * W;
  {
    LaunchScope ls(st, label, 0.0, (mask ? 12.0 : 8.0) * n);
    if (mask) p8_pack_kernel<true><<<grid, PK_THREADS, 0, st>>>(x, mask, (__nv_bfloat16*)p8, acc, B, C, H, W);
    else p8_pack_kernel<false><<<grid, PK_THREADS, 0, st>>>(x, mask, (__nv_bfloat16*)p8, acc, B, C, H, W);
  }
  return (int)cudaGetLastError();
}

extern "C" int nconv_b200_p8_pack(const float* x, void* p8, int B, int C, int H, int W, void* stream) {
  if (!x || !p8 || B <= 0 || C <= 0 || H <= 0 || W <= 0) return NCONV_ERR_BAD_ARG;
  return p8_pack_launch(x, nullptr, p8, nullptr, B, C, H, W, "p8_pack", (cudaStream_t)stream);
}

namespace nconv { __global__ void p8_sums_out_kernel(const double* __restrict__ a, float* __restrict__ o, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) o[i] = (float)a[i];
} }

// ReLU(conv + bias) backward straight into the tensor-core layout: p8 = P8(gy * [y > 0]) (y null: no mask), gbias[c] = the
// per-channel sum of the masked gradient (null: not wanted).  ws: C doubles of scratch.  Replaces relu_bias_bwd + p8_pack
// where both consumers of the masked gradient (data and weight gradient) run on tcgen05 (train_step2.py:62-64).
extern "C" int nconv_b200_relu_bias_bwd_p8(const float* gy, const float* y, void* p8, float* gbias, void* ws, int B, int C, int H,
                                           int W, void* stream) {
  if (!gy || !p8 || B <= 0 || C <= 0 || H <= 0 || W <= 0 || (gbias && !ws)) return NCONV_ERR_BAD_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  if (gbias) NCONV_RETURN_IF_ERR(cudaMemsetAsync(ws, 0, sizeof(double) * C, st));
  const int rc = p8_pack_launch(gy, y, p8, gbias ? (double*)ws : nullptr, B, C, H, W, "relu_bias_bwd_p8", st);
  if (rc) return rc;
  if (gbias) {
    LaunchScope ls(st, "channel_sum_out", 0.0, 0.0);
    p8_sums_out_kernel<<<(C + 127) / 128, 128, 0, st>>>((const double*)ws, gbias, C);
  }
  return (int)cudaGetLastError();
}

extern "C" int nconv_b200_p8_unpack_into(const void* p8, float* x, int B, int C, int H, int W, int Ctot, int c0, void* stream) {
  if (!x || !p8 || B <= 0 || C <= 0 || H <= 0 || W <= 0 || c0 < 0 || c0 + C > Ctot) return NCONV_ERR_BAD_ARG;
  const P8Geom g = P8Geom::make(B, C, H, W);
  if ((long long)B * g.C8 > 65535) return NCONV_ERR_UNSUPPORTED;
  dim3 grid((W + PK_THREADS - 1) / PK_THREADS, (H + PK_ROWS - 1) / PK_ROWS, B * g.C8);
  {
    LaunchScope ls((cudaStream_t)stream, "p8_unpack", 0.0, 8.0 * B * C * H * W);
    p8_unpack_kernel<<<grid, PK_THREADS, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)p8, x, B, C, H, W, Ctot, c0);
  }
  return (int)cudaGetLastError();
}

extern "C" int nconv_b200_p8_pack_ex(const float* x, void* p8, int B, int C, int H, int W, int Ctot, int c0, int phases, void* stream) {
  if (!x || !p8 || B <= 0 || C <= 0 || H <= 0 || W <= 0 || c0 < 0 || c0 + C > Ctot) return NCONV_ERR_BAD_ARG;
  if (phases && (C % 8)) return NCONV_ERR_UNSUPPORTED;
  const P8Geom g = P8Geom::make(B, phases ? 4 * C : C, H, W);
  if ((long long)B * g.C8 > 65535) return NCONV_ERR_UNSUPPORTED;
  dim3 grid((g.Wp + PK_THREADS - 1) / PK_THREADS, (g.Hp + PK_ROWS - 1) / PK_ROWS, B * g.C8);
  cudaStream_t st = (cudaStream_t)stream;
  {
    LaunchScope ls(st, phases ? "p8_pack_phases" : "p8_pack", 0.0, 8.0 * B * (phases ? 4.0 : 1.0) * C * H * W);
    if (phases) p8_pack_ex_kernel<true><<<grid, PK_THREADS, 0, st>>>(x, (__nv_bfloat16*)p8, B, C, H, W, Ctot, c0);
    else p8_pack_ex_kernel<false><<<grid, PK_THREADS, 0, st>>>(x, (__nv_bfloat16*)p8, B, C, H, W, Ctot, c0);
  }
  return (int)cudaGetLastError();
}

extern "C" int nconv_b200_convS2_phase_weights(const float* w, float* w3d, int CO, int CI, void* stream) {
  if (!w || !w3d || CI <= 0 || CO <= 0) return NCONV_ERR_BAD_ARG;
  const int total = 4 * CI * CO * 9;
  {
    LaunchScope ls((cudaStream_t)stream, "convS2_phase_weights", 0.0, 4.0 * (total + CI * CO * 9));
    convS2_phase_weights_kernel<<<(total + 255) / 256, 256, 0, (cudaStream_t)stream>>>(w, w3d, CO, CI);
  }
  return (int)cudaGetLastError();
}

extern "C" int nconv_b200_convS2_phase_wgrad(const float* gw3, float* gw, int CO, int CI, void* stream) {
  if (!gw3 || !gw || CI <= 0 || CO <= 0) return NCONV_ERR_BAD_ARG;
  const int total = CO * CI * 9;
  {
    LaunchScope ls((cudaStream_t)stream, "convS2_phase_wgrad", 0.0, 8.0 * total);
    convS2_phase_wgrad_kernel<<<(total + 255) / 256, 256, 0, (cudaStream_t)stream>>>(gw3, gw, CO, CI);
  }
  return (int)cudaGetLastError();
}

extern "C" int nconv_b200_convT_phase_weights(const float* w, float* w3, int CI, int CO, void* stream) {
  if (!w || !w3 || CI <= 0 || CO <= 0) return NCONV_ERR_BAD_ARG;
  const int total = CI * 4 * CO * 9;
  {
    LaunchScope ls((cudaStream_t)stream, "convT_phase_weights", 0.0, 4.0 * (total + CI * CO * 16));
    convT_phase_weights_kernel<<<(total + 255) / 256, 256, 0, (cudaStream_t)stream>>>(w, w3, CI, CO);
  }
  return (int)cudaGetLastError();
}

extern "C" int nconv_b200_convT_phase_wgrad(const float* gw3, float* gw, int CIF, int CO, int lead, void* stream) {
  if (!gw3 || !gw || CIF <= 0 || CO <= 0 || lead < 0) return NCONV_ERR_BAD_ARG;
  const int total = CIF * CO * 16;
  {
    LaunchScope ls((cudaStream_t)stream, "convT_phase_wgrad", 0.0, 8.0 * total);
    convT_phase_wgrad_kernel<<<(total + 255) / 256, 256, 0, (cudaStream_t)stream>>>(gw3, gw, CIF, CO, lead);
  }
  return (int)cudaGetLastError();
}

extern "C" int nconv_b200_p8_unpack(const void* p8, float* x, int B, int C, int H, int W, void* stream) {
  return nconv_b200_p8_unpack_into(p8, x, B, C, H, W, C, 0, stream);
}

extern "C" int nconv_b200_p8_zero_ring(void* p8, int B, int C, int H, int W, void* stream) {
  if (!p8 || B <= 0 || C <= 0 || H <= 0 || W <= 0) return NCONV_ERR_BAD_ARG;
  const P8Geom g = P8Geom::make(B, C, H, W);
  const int planes = B * g.C8 * 2;
  const size_t total = (size_t)planes * (2 * g.Wp + 2 * (g.Hp - 2));
  int blocks = (int)((total + 255) / 256); if (blocks > 148 * 8) blocks = 148 * 8;
  {
    LaunchScope ls((cudaStream_t)stream, "p8_zero_ring", 0.0, 16.0 * total);
    p8_zero_ring_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>((__nv_bfloat16*)p8, planes, g.Hp, g.Wp);
  }
  return (int)cudaGetLastError();
}

namespace nconv {
// zero row 0 and column 0 of every P8Q phase plane
__global__ void p8q_zero_ring_kernel(__nv_bfloat16* __restrict__ p, size_t planes, int Hq1, int Wq1) {
  const int per = Wq1 + Hq1 - 1;
  const size_t total = planes * per;
  const uint4 z = make_uint4(0, 0, 0, 0);
  for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
    int i = (int)(idx % per);
    size_t pl = idx / per;
    int y = (i < Wq1) ? 0 : i - Wq1 + 1, x = (i < Wq1) ? i : 0;
    *reinterpret_cast<uint4*>(p + (pl * Hq1 * Wq1 + (size_t)y * Wq1 + x) * 8) = z;
  }
}
}  // namespace nconv

extern "C" size_t nconv_b200_p8q_bytes(int B, int C, int H, int W) {
  if (B <= 0 || C <= 0 || H <= 0 || W <= 0 || ((H | W) & 1)) return 0;
  return P8QGeom::make(B, C, H, W).total_elems() * sizeof(__nv_bfloat16) + P8_TAIL_SLACK_BYTES;
}

extern "C" int nconv_b200_p8q_zero_ring(void* p8q, int B, int C, int H, int W, void* stream) {
  if (!p8q || B <= 0 || C <= 0 || H <= 0 || W <= 0 || ((H | W) & 1)) return NCONV_ERR_BAD_ARG;
  const P8QGeom g = P8QGeom::make(B, C, H, W);
  const size_t planes = (size_t)B * g.C8 * 2 * 4;
  const size_t total = planes * (g.Wq1 + g.Hq1 - 1);
  int blocks = (int)((total + 255) / 256); if (blocks > 148 * 8) blocks = 148 * 8;
  {
    LaunchScope ls((cudaStream_t)stream, "p8q_zero_ring", 0.0, 16.0 * total);
    p8q_zero_ring_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>((__nv_bfloat16*)p8q, planes, g.Hq1, g.Wq1);
  }
  return (int)cudaGetLastError();
}

namespace nconv {
__global__ void p8q_pack_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ p, int B, int C, int H, int W) {
  const P8QGeom g = P8QGeom::make(B, C, H, W);
  const size_t total = (size_t)B * g.C8 * 4 * g.Hq1 * g.Wq1;
  for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
    int xq = (int)(idx % g.Wq1);
    size_t r = idx / g.Wq1;
    int yq = (int)(r % g.Hq1); r /= g.Hq1;
    int ph = (int)(r & 3); r >>= 2;
    int ch = (int)(r % g.C8);
    int n = (int)(r / g.C8);
    const int yy = 2 * (yq - 1) + (ph >> 1), xx = 2 * (xq - 1) + (ph & 1);
    const bool inside = yq >= 1 && xq >= 1;
    float f[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int c = ch * 8 + e;
      f[e] = (inside && c < C) ? __ldg(x + (((size_t)n * C + c) * H + yy) * W + xx) : 0.f;
    }
    uint4 hi, lo;
    split8(f, hi, lo);
    const size_t off = (size_t)ph * g.phase_elems() + ((size_t)yq * g.Wq1 + xq) * 8;
    *reinterpret_cast<uint4*>(p + g.slab(n, ch, 0) + off) = hi;
    *reinterpret_cast<uint4*>(p + g.slab(n, ch, 1) + off) = lo;
  }
}
}  // namespace nconv

extern "C" int nconv_b200_p8q_pack(const float* x, void* p8q, int B, int C, int H, int W, void* stream) {
  if (!x || !p8q || B <= 0 || C <= 0 || H <= 0 || W <= 0 || ((H | W) & 1)) return NCONV_ERR_BAD_ARG;
  const P8QGeom g = P8QGeom::make(B, C, H, W);
  const size_t total = (size_t)B * g.C8 * 4 * g.Hq1 * g.Wq1;
  int blocks = (int)((total + 255) / 256); if (blocks > 148 * 16) blocks = 148 * 16;
  {
    LaunchScope ls((cudaStream_t)stream, "p8q_pack", 0.0, 8.0 * B * C * H * W);
    p8q_pack_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(x, (__nv_bfloat16*)p8q, B, C, H, W);
  }
  return (int)cudaGetLastError();
}
