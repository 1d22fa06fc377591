// fp32 SIMT direct convolutions for the RGB-guided stack of SETP2
// (/root/reference/models/step2.py:130-297): KxK (K in {1,3}), stride 1/2, two-source channel
// concat on the load side, per-channel scale/shift (folded eval BatchNorm or bias), ReLU and a
// residual add in the epilogue; and the ConvTranspose2d(4,2,1) of Basic2dTrans as four sub-pixel
// 2x2 convolutions.  These are the exact-fp32 path: every layer runs here first; layers with
// Ci,Co >= 32 are the tensor-core candidates.
#include "common.cuh"
#include "nconv_b200.h"
#include "launch_count.h"
#include <cstdio>
#include <type_traits>

namespace nconv {

constexpr int CT_TW = 64;        // output tile width
constexpr int CT_PX = 4;         // pixels per thread
constexpr int CT_THREADS = 256;

struct ConvArgs {
  const float* a; const float* b; const float* w; const float* scale; const float* shift;
  const float* residual; float* y;
  int B, CA, CB, CO, H, W, HO, WO, relu, vec_ok;
};

template <int K, int S, int CO_T>
struct ConvCfg {
  static constexpr int TH = (S == 1) ? 16 : 8;               // output rows per CTA
  static constexpr int ROWS_PER_PASS = CT_THREADS / (CT_TW / CT_PX);  // 16 thread rows
  static constexpr int PAD = K / 2;
  static constexpr int OFF = (4 - (PAD & 3)) & 3;
  static constexpr int WIN = (CT_PX - 1) * S + K;            // input cols per thread
  static constexpr int NV = (OFF + WIN + 3) / 4;
  static constexpr int SW = CT_TW * S + 8;                   // staged row width
  static constexpr int RH = (TH - 1) * S + K;                // staged rows
  static constexpr int CCH = (S == 1) ? 8 : 4;               // channels per stage
  static constexpr int W_FLOATS = CCH * K * K * CO_T;
  static constexpr int TILE_FLOATS = CCH * RH * SW;
  static constexpr int BUF_FLOATS = W_FLOATS + TILE_FLOATS;
  static constexpr size_t BYTES = sizeof(float) * 2 * BUF_FLOATS;   // double-buffered chunks
};

// Each CTA: TH x 64 output pixels x CO_T output channels of one image.  With S==2 only 8 of the
// 16 thread rows would have work, so the thread block is reshaped to 8 rows x 16 strips x 2
// channel halves (each thread then owns CO_T/2 channels).
template <int K, int S, int CO_T>
__global__ void __launch_bounds__(CT_THREADS, 2) conv_fwd_tiled(const ConvArgs a) {
  using C = ConvCfg<K, S, CO_T>;
  constexpr int TH = C::TH, PAD = C::PAD, OFF = C::OFF, NV = C::NV, SW = C::SW, RH = C::RH,
                CCH = C::CCH;
  constexpr int SPLIT = (S == 1) ? 1 : 2;        // channel halves per CTA
  constexpr int CO_TT = CO_T / SPLIT;            // channels per thread
  constexpr int WV = (CO_TT % 4 == 0) ? 4 : 1;
  static_assert(CO_T % SPLIT == 0, "");
  static_assert(4 * NV + 4 * S * (CT_TW / CT_PX - 1) <= SW, "staged row too narrow");

  extern __shared__ __align__(16) float smem[];   // 2 x { weights [CCH][K][K][CO_T] | inputs [CCH][RH][SW] }

  const int tid = threadIdx.x;
  const int tx = tid & 15;
  const int ty = (tid >> 4) % TH;
  const int half = (tid >> 4) / TH;  // 0 when SPLIT==1
  const int NG = a.CO / CO_T;
  const int b = blockIdx.z / NG, cog = blockIdx.z % NG;
  const int ox0 = blockIdx.x * CT_TW, oy0 = blockIdx.y * TH;
  const int CI = a.CA + a.CB;

  // Even channel counts run the FMA stream packed (FFMA2, common.cuh): the weight PAIR of two adjacent output
  // channels is a 64-bit half of the broadcast LDS.128, the activation the scalar-broadcast operand; same
  // products in the same order as the scalar loop, two per issue slot.
  constexpr bool PACK = (CO_TT % 4 == 0);
  constexpr int CO_P = PACK ? CO_TT / 2 : 1;
  float acc[PACK ? 1 : CO_TT][CT_PX];
  u64 acc2[CO_P][CT_PX];
#pragma unroll
  for (int o = 0; o < (PACK ? 1 : CO_TT); ++o)
#pragma unroll
    for (int p = 0; p < CT_PX; ++p) acc[o][p] = 0.f;
#pragma unroll
  for (int o = 0; o < CO_P; ++o)
#pragma unroll
    for (int p = 0; p < CT_PX; ++p) acc2[o][p] = 0ull;

  const int gx0a = ox0 * S - PAD - OFF;
  const int gy0 = oy0 * S - PAD;

  // Stage the chunk of channels [c0, c0+CCH) into one buffer.  ASYNC (128-bit rows, a.vec_ok: a quad is entirely
  // inside or outside the image): cp.async copies, zero-filled outside, issued one chunk ahead of the FFMA2 loop.
  // (ncu before: 45 % of the warp samples were the STS of the staging loop waiting for its own global load — ten
  // sequential DRAM round trips per chunk against 2 300 FFMA2.)  Otherwise: load + store as before.
  auto stage = [&](int c0, float* buf, auto async_tag) {
    constexpr bool ASYNC = decltype(async_tag)::value;
    float* b_w = buf;
    float* b_in = buf + C::W_FLOATS;
    const int nch = min(CCH, CI - c0);
    for (int idx = tid; idx < nch * K * K * CO_T; idx += CT_THREADS) {
      int o = idx % CO_T;
      int r = idx / CO_T;
      int v = r % K; r /= K;
      int u = r % K; r /= K;
      const float* src = &a.w[(((size_t)(cog * CO_T + o) * CI + c0 + r) * K + u) * K + v];
      if constexpr (ASYNC) cp_async4(b_w + idx, src, true); else b_w[idx] = __ldg(src);
    }
    const int slots = nch * RH * (SW / 4);
    for (int idx = tid; idx < slots; idx += CT_THREADS) {
      int j = idx % (SW / 4);
      int r = (idx / (SW / 4)) % RH;
      int cc = idx / ((SW / 4) * RH);
      int cs = c0 + cc;
      int gy = gy0 + r, gx = gx0a + 4 * j;
      const bool row_ok = gy >= 0 && gy < a.H;
      const float* src = a.a;
      if (row_ok)
        src = (cs < a.CA) ? a.a + ((size_t)(b * a.CA + cs) * a.H + gy) * a.W
                          : a.b + ((size_t)(b * a.CB + (cs - a.CA)) * a.H + gy) * a.W;
      float* dst = b_in + (cc * RH + r) * SW + 4 * j;
      if constexpr (ASYNC) {
        const bool ok = row_ok && gx >= 0 && gx + 3 < a.W;
        cp_async16(dst, ok ? src + gx : a.a, ok);
      } else {
        float4 val = make_float4(0.f, 0.f, 0.f, 0.f);
        if (row_ok) {
          if (a.vec_ok && gx >= 0 && gx + 3 < a.W) {
            val = __ldg(reinterpret_cast<const float4*>(src + gx));
          } else {
            float t[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int e = 0; e < 4; ++e)
              if (gx + e >= 0 && gx + e < a.W) t[e] = __ldg(src + gx + e);
            val = make_float4(t[0], t[1], t[2], t[3]);
          }
        }
        *reinterpret_cast<float4*>(dst) = val;
      }
    }
  };

  const bool async = a.vec_ok != 0;
  if (async) stage(0, smem, std::true_type{});
  int it = 0;
#pragma unroll 1
  for (int c0 = 0; c0 < CI; c0 += CCH, ++it) {
    const int nch = min(CCH, CI - c0);
    float* buf = smem + (it & 1) * C::BUF_FLOATS;
    if (async) {
      cp_async_wait_all();
      __syncthreads();               // chunk visible to all; everyone is done with the other buffer
      if (c0 + CCH < CI) stage(c0 + CCH, smem + ((it & 1) ^ 1) * C::BUF_FLOATS, std::true_type{});
    } else {
      __syncthreads();
      stage(c0, buf, std::false_type{});
      __syncthreads();
    }
    const float* s_w = buf;
    const float* s_in = buf + C::W_FLOATS;

#pragma unroll 1
    for (int cc = 0; cc < nch; ++cc) {
      const float* wrow = s_w + cc * K * K * CO_T + half * CO_TT;
#pragma unroll
      for (int u = 0; u < K; ++u) {
        float xr[4 * NV];
        const float* srow = s_in + (cc * RH + ty * S + u) * SW + 4 * S * tx;
#pragma unroll
        for (int q = 0; q < NV; ++q) {
          float4 t = *reinterpret_cast<const float4*>(srow + 4 * q);
          xr[4 * q] = t.x; xr[4 * q + 1] = t.y; xr[4 * q + 2] = t.z; xr[4 * q + 3] = t.w;
        }
#pragma unroll
        for (int v = 0; v < K; ++v) {
          if constexpr (PACK) {
#pragma unroll
            for (int o = 0; o < CO_P; o += 2) {
              const ulonglong2 t = *reinterpret_cast<const ulonglong2*>(wrow + (u * K + v) * CO_T + 2 * o);
#pragma unroll
              for (int p = 0; p < CT_PX; ++p) {
                ffma2(acc2[o][p], t.x, xr[OFF + p * S + v]);
                ffma2(acc2[o + 1][p], t.y, xr[OFF + p * S + v]);
              }
            }
          } else {
          float wv[CO_TT];
#pragma unroll
          for (int o = 0; o < CO_TT; o += WV) {
            if (WV == 4) {
              float4 t = *reinterpret_cast<const float4*>(wrow + (u * K + v) * CO_T + o);
              wv[o] = t.x; wv[o + 1] = t.y; wv[o + 2] = t.z; wv[o + 3] = t.w;
            } else {
              wv[o] = wrow[(u * K + v) * CO_T + o];
            }
          }
#pragma unroll
          for (int o = 0; o < CO_TT; ++o)
#pragma unroll
            for (int p = 0; p < CT_PX; ++p)
              acc[o][p] = fmaf(wv[o], xr[OFF + p * S + v], acc[o][p]);
          }
        }
      }
    }
  }

  const int oy = oy0 + ty, ox = ox0 + CT_PX * tx;
  if (oy >= a.HO) return;
#pragma unroll
  for (int o = 0; o < CO_TT; ++o) {
    const int oc = cog * CO_T + half * CO_TT + o;
    const float sc = a.scale ? __ldg(a.scale + oc) : 1.f;
    const float sh = a.shift ? __ldg(a.shift + oc) : 0.f;
    size_t base = (((size_t)b * a.CO + oc) * a.HO + oy) * a.WO + ox;
    float r[CT_PX];
#pragma unroll
    for (int p = 0; p < CT_PX; ++p) {
      const float av = PACK ? ((o & 1) ? hi32f(acc2[PACK ? o >> 1 : 0][p]) : lo32f(acc2[PACK ? o >> 1 : 0][p])) : acc[PACK ? 0 : o][p];
      float v = fmaf(av, sc, sh);
      if (a.relu) v = fmaxf(v, 0.f);
      r[p] = v;
    }
    if (a.vec_ok && (a.WO & 3) == 0 && ox + 3 < a.WO) {
      if (a.residual) {
        float4 q = __ldg(reinterpret_cast<const float4*>(a.residual + base));
        r[0] += q.x; r[1] += q.y; r[2] += q.z; r[3] += q.w;
      }
      *reinterpret_cast<float4*>(a.y + base) = make_float4(r[0], r[1], r[2], r[3]);
    } else {
#pragma unroll
      for (int p = 0; p < CT_PX; ++p)
        if (ox + p < a.WO) a.y[base + p] = r[p] + (a.residual ? __ldg(a.residual + base + p) : 0.f);
    }
  }
}

template <int K, int S, int CO_T>
static int launch_conv(const ConvArgs& a, cudaStream_t st) {
  using C = ConvCfg<K, S, CO_T>;
  auto kern = conv_fwd_tiled<K, S, CO_T>;
  static bool attr_done[NCONV_MAX_DEVICES] = {};
  if (C::BYTES > 48 * 1024) {
    int dev = 0;
    NCONV_RETURN_IF_ERR(cudaGetDevice(&dev));
    if (dev >= NCONV_MAX_DEVICES || !attr_done[dev]) {
      NCONV_RETURN_IF_ERR(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::BYTES));
      if (dev < NCONV_MAX_DEVICES) attr_done[dev] = true;
    }
  }
  dim3 grid(ceil_div(a.WO, CT_TW), ceil_div(a.HO, C::TH), a.B * (a.CO / CO_T));
  const int CI = a.CA + a.CB;
  const double pin = (double)a.B * a.H * a.W, pout = (double)a.B * a.HO * a.WO;
  char label[64];
  snprintf(label, sizeof(label), "conv%dx%d s%d %d->%d @%dx%d", K, K, S, CI, a.CO, a.H, a.W);
  {
    LaunchScope ls(st, label, 2.0 * CI * a.CO * K * K * pout,
                   4.0 * (CI * pin + a.CO * pout * (a.residual ? 2.0 : 1.0)));
    kern<<<grid, CT_THREADS, C::BYTES, st>>>(a);
  }
  return (int)cudaGetLastError();
}

// ---- ConvTranspose2d(k=4, s=2, p=1) as four sub-pixel 2x2 convolutions ---------------------
// out[o, 2a+py, 2b+px] = sum_i sum_{t,s in {0,1}} in[i, a+py-t, b+px-s] * W[i,o,(1-py)+2t,(1-px)+2s]
// A thread owns input positions (a, b0) and (a, b0+1) -> a 2x4 patch of outputs, for 8 channels.
constexpr int TT_CO = 8, TT_CCH = 8, TT_TH = 16, TT_TWP = 32, TT_SW = TT_TWP + 4;
struct ConvTArgs {
  const float* a; const float* b; const float* w; const float* scale; const float* shift; float* y;
  int B, CA, CB, CO, H, W, relu;
};

constexpr int TT_W_FLOATS = TT_CCH * 16 * TT_CO, TT_IN_FLOATS = TT_CCH * (TT_TH + 2) * TT_SW;
constexpr int TT_BUF_FLOATS = TT_W_FLOATS + TT_IN_FLOATS;
constexpr size_t TT_SMEM = sizeof(float) * 2 * TT_BUF_FLOATS;      // double-buffered chunks

__global__ void __launch_bounds__(256, 2) convT4x4s2_kernel(const ConvTArgs a) {
  extern __shared__ __align__(16) float tt_smem[];   // 2 x { weights [CCH][16][CO] | inputs [CCH][TH+2][SW] }
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  const int NG = a.CO / TT_CO;
  const int b = blockIdx.z / NG, cog = blockIdx.z % NG;
  const int a0 = blockIdx.y * TT_TH, b0 = blockIdx.x * TT_TWP;
  const int CI = a.CA + a.CB;

  float acc[TT_CO][2][4];  // [co][py][out col 0..3]
#pragma unroll
  for (int o = 0; o < TT_CO; ++o)
#pragma unroll
    for (int r = 0; r < 2; ++r)
#pragma unroll
      for (int q = 0; q < 4; ++q) acc[o][r][q] = 0.f;

  // the chunk of channels [c0, c0+CCH) is copied global -> shared with cp.async (4-byte copies: the halo starts at
  // column b0-1; zero-filled outside the image) one chunk ahead of the FMA loop
  auto stage = [&](int c0, float* buf) {
    float* b_w = buf;
    float* b_in = buf + TT_W_FLOATS;
    const int nch = min(TT_CCH, CI - c0);
    for (int idx = tid; idx < nch * 16 * TT_CO; idx += 256) {
      int o = idx % TT_CO;
      int tap = (idx / TT_CO) % 16;
      int cc = idx / (TT_CO * 16);
      cp_async4(b_w + idx, &a.w[((size_t)(c0 + cc) * a.CO + cog * TT_CO + o) * 16 + tap], true);
    }
    for (int idx = tid; idx < nch * (TT_TH + 2) * TT_SW; idx += 256) {
      int j = idx % TT_SW;
      int r = (idx / TT_SW) % (TT_TH + 2);
      int cc = idx / (TT_SW * (TT_TH + 2));
      int cs = c0 + cc;
      int gy = a0 - 1 + r, gx = b0 - 1 + j;
      const bool ok = gy >= 0 && gy < a.H && gx >= 0 && gx < a.W;
      const float* src = a.a;
      if (ok)
        src = (cs < a.CA) ? a.a + ((size_t)(b * a.CA + cs) * a.H + gy) * a.W + gx
                          : a.b + ((size_t)(b * a.CB + cs - a.CA) * a.H + gy) * a.W + gx;
      cp_async4(b_in + idx, src, ok);                   // [cc][r][j] is idx in memory order
    }
  };

  stage(0, tt_smem);
  int it = 0;
#pragma unroll 1
  for (int c0 = 0; c0 < CI; c0 += TT_CCH, ++it) {
    const int nch = min(TT_CCH, CI - c0);
    float* buf = tt_smem + (it & 1) * TT_BUF_FLOATS;
    cp_async_wait_all();
    __syncthreads();                 // chunk visible to all; everyone is done with the other buffer
    if (c0 + TT_CCH < CI) stage(c0 + TT_CCH, tt_smem + ((it & 1) ^ 1) * TT_BUF_FLOATS);
    const float* s_w = buf;
    const float (*s_in)[TT_TH + 2][TT_SW] = reinterpret_cast<const float (*)[TT_TH + 2][TT_SW]>(buf + TT_W_FLOATS);
#pragma unroll 1
    for (int cc = 0; cc < nch; ++cc) {
      float in[3][4];  // rows a-1..a+1, cols b-1..b+2 (b = b0 + 2*tx)
#pragma unroll
      for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int q = 0; q < 4; ++q) in[r][q] = s_in[cc][ty + r][2 * tx + q];
      const float* wc = s_w + cc * 16 * TT_CO;
#pragma unroll
      for (int kh = 0; kh < 4; ++kh) {
        // kh = (1-py) + 2t  ->  py = 1 - (kh & 1), t = kh >> 1, input row offset dr = py - t
        const int py = 1 - (kh & 1), dr = py - (kh >> 1);
#pragma unroll
        for (int kw = 0; kw < 4; ++kw) {
          const int px = 1 - (kw & 1), dc = px - (kw >> 1);
          float wv[TT_CO];
#pragma unroll
          for (int o = 0; o < TT_CO; o += 4) {
            float4 t = *reinterpret_cast<const float4*>(wc + (kh * 4 + kw) * TT_CO + o);
            wv[o] = t.x; wv[o + 1] = t.y; wv[o + 2] = t.z; wv[o + 3] = t.w;
          }
#pragma unroll
          for (int pos = 0; pos < 2; ++pos) {
            const float x = in[1 + dr][1 + pos + dc];
#pragma unroll
            for (int o = 0; o < TT_CO; ++o) acc[o][py][2 * pos + px] = fmaf(wv[o], x, acc[o][py][2 * pos + px]);
          }
        }
      }
    }
  }
  const int ia = a0 + ty, ib = b0 + 2 * tx;
  if (ia >= a.H) return;
  const int HO = 2 * a.H, WO = 2 * a.W;
#pragma unroll
  for (int o = 0; o < TT_CO; ++o) {
    const int oc = cog * TT_CO + o;
    const float sc = a.scale ? __ldg(a.scale + oc) : 1.f;
    const float sh = a.shift ? __ldg(a.shift + oc) : 0.f;
#pragma unroll
    for (int py = 0; py < 2; ++py) {
      size_t base = (((size_t)b * a.CO + oc) * HO + 2 * ia + py) * WO + 2 * ib;
      float r[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        float v = fmaf(acc[o][py][q], sc, sh);
        r[q] = a.relu ? fmaxf(v, 0.f) : v;
      }
      if (ib + 1 < a.W && (WO & 3) == 0 && ((uintptr_t)(a.y) & 15) == 0) {
        *reinterpret_cast<float4*>(a.y + base) = make_float4(r[0], r[1], r[2], r[3]);
      } else {
#pragma unroll
        for (int q = 0; q < 4; ++q)
          if (2 * ib + q < WO) a.y[base + q] = r[q];
      }
    }
  }
}

}  // namespace nconv

using namespace nconv;

extern "C" int nconv_b200_conv2d_fwd(const nconv_b200_conv_desc* d, void* stream) {
  if (!d || !d->a || !d->w || !d->y) return NCONV_ERR_BAD_ARG;
  if (d->B <= 0 || d->CA <= 0 || d->CB < 0 || d->CO <= 0 || d->H <= 0 || d->W <= 0) return NCONV_ERR_BAD_ARG;
  if (d->CB > 0 && !d->b) return NCONV_ERR_BAD_ARG;
  if (!(d->K == 1 || d->K == 3) || !(d->stride == 1 || d->stride == 2)) return NCONV_ERR_UNSUPPORTED;
  ConvArgs a;
  a.a = d->a; a.b = d->b; a.w = d->w; a.scale = d->scale; a.shift = d->shift;
  a.residual = d->residual; a.y = d->y;
  a.B = d->B; a.CA = d->CA; a.CB = d->CB; a.CO = d->CO; a.H = d->H; a.W = d->W; a.relu = d->relu;
  const int pad = d->K / 2;
  a.HO = (d->H + 2 * pad - d->K) / d->stride + 1;
  a.WO = (d->W + 2 * pad - d->K) / d->stride + 1;
  a.vec_ok = (d->W % 4 == 0) && aligned16(d->a) && (d->CB == 0 || aligned16(d->b)) &&
             aligned16(d->y) && (!d->residual || aligned16(d->residual));
  cudaStream_t st = (cudaStream_t)stream;
  const bool co16 = d->CO % 16 == 0;
  if (d->K == 3 && d->stride == 1) return co16 ? launch_conv<3, 1, 16>(a, st) : launch_conv<3, 1, 1>(a, st);
  if (d->K == 3 && d->stride == 2) return co16 ? launch_conv<3, 2, 16>(a, st) : NCONV_ERR_UNSUPPORTED;
  if (d->K == 1 && d->stride == 1) return co16 ? launch_conv<1, 1, 16>(a, st) : launch_conv<1, 1, 1>(a, st);
  if (d->K == 1 && d->stride == 2) return co16 ? launch_conv<1, 2, 16>(a, st) : NCONV_ERR_UNSUPPORTED;
  return NCONV_ERR_UNSUPPORTED;
}

extern "C" int nconv_b200_convT4x4s2_fwd(const float* a, const float* b, const float* w,
                                         const float* scale, const float* shift, float* y, int B,
                                         int CA, int CB, int CO, int H, int W, int relu,
                                         void* stream) {
  if (!a || !w || !y || B <= 0 || CA <= 0 || CB < 0 || CO <= 0 || H <= 0 || W <= 0) return NCONV_ERR_BAD_ARG;
  if (CB > 0 && !b) return NCONV_ERR_BAD_ARG;
  if (CO % TT_CO != 0) return NCONV_ERR_UNSUPPORTED;
  ConvTArgs t;
  t.a = a; t.b = b; t.w = w; t.scale = scale; t.shift = shift; t.y = y;
  t.B = B; t.CA = CA; t.CB = CB; t.CO = CO; t.H = H; t.W = W; t.relu = relu;
  dim3 grid(ceil_div(W, TT_TWP), ceil_div(H, TT_TH), B * (CO / TT_CO));
  char label[64];
  snprintf(label, sizeof(label), "convT4x4s2 %d->%d @%dx%d", CA + CB, CO, H, W);
  {
    const double pin = (double)B * H * W;
    LaunchScope ls((cudaStream_t)stream, label, 2.0 * (CA + CB) * CO * 16 * pin,
                   4.0 * ((CA + CB) * pin + 4.0 * CO * pin));
    static bool attr_done[NCONV_MAX_DEVICES] = {};
    int dev = 0;
    NCONV_RETURN_IF_ERR(cudaGetDevice(&dev));
    if (dev >= NCONV_MAX_DEVICES || !attr_done[dev]) {
      NCONV_RETURN_IF_ERR(cudaFuncSetAttribute(convT4x4s2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TT_SMEM));
      if (dev < NCONV_MAX_DEVICES) attr_done[dev] = true;
    }
    convT4x4s2_kernel<<<grid, 256, TT_SMEM, (cudaStream_t)stream>>>(t);
  }
  return (int)cudaGetLastError();
}
