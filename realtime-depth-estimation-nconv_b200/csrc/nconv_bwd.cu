// NConv2d backward (derived from /root/reference/models/step1.py:116-149, which
// the reference differentiates with autograd through ~9 eager ops).
//
//   s[o]  = sum W[o]            D = cout*s      N/(D+eps) = y - b        (from saved outputs)
//   gN = gy/(D+eps)             gD = gcout/s - gN*(y-b)
//   gx = convT(gN,W) * c        gc = convT(gN,W) * x + convT(gD,W)
//   gW[o,i,u,v] = sum_{b,p} gN[o,p]*(x*c)[i,p+(u,v)-pad] + gD[o,p]*c[i,p+(u,v)-pad]
//                 - (sum_{b,p} gcout[o,p]*cout[o,p]) / s[o]
//   gb[o] = sum_{b,p} gy[o,p]
//
// Three kernels: (1) per-channel weight sums; (2) data gradient = the dual-stream tiled
// correlation of nconv_tiled.cuh run on (gN,gD) with flipped weights; (3) weight gradient as a
// persistent tiled reduction with per-CTA partials and a fixed-order (deterministic) final sum.
#include "nconv_tiled.cuh"
#include "nconv_bwd_stream.cuh"
#include <cstdio>

namespace nconv {

__global__ void wsum_kernel(const float* __restrict__ w, float* __restrict__ s, int Co, int n) {
  int o = blockIdx.x;
  float acc = 0.f;
  for (int j = threadIdx.x; j < n; j += 32) acc += __ldg(w + (size_t)o * n + j);
  acc = warp_sum(acc);
  if (threadIdx.x == 0) s[o] = acc;
}

// ---- weight gradient ------------------------------------------------------------------------
constexpr int WG_TH = 8;      // output-space tile rows
constexpr int WG_TW = 64;
constexpr int WG_THREADS = 256;

template <int CI, int CO, int K>
struct WGradCfg {
  static constexpr int CCH = CI < 4 ? CI : 4;         // input channels per CTA
  static constexpr int CO_T = CO < 8 ? CO : 8;        // output channels per CTA
  static constexpr int NCI = CI / CCH, NCO = CO / CO_T;
  static constexpr int RH = WG_TH + K - 1;
  static constexpr int P = CCH * K;                   // (i,u) pairs
  static constexpr int NG = WG_THREADS / P;           // pixel-strip groups
  static constexpr int G_FLOATS = CO_T * WG_TH * WG_TW;       // one stream of gN or gD
  static constexpr int X_FLOATS = CCH * RH * SW;
  static constexpr int RED_FLOATS = P * CO_T * K;
  static constexpr int PART_FLOATS = 8 * 2 * 8;               // [warp][gb|T][o]
  static constexpr size_t BYTES = sizeof(float) * (2 * G_FLOATS + 2 * X_FLOATS + RED_FLOATS + PART_FLOATS);
  static constexpr int OUT_PER_CTA = CO_T * CCH * K * K;      // weight partials per CTA
};

struct WGradArgs {
  const float* gy; const float* gco; const float* y; const float* co;
  const float* x; const float* c; const float* bias; const float* wsum;
  float* partial;     // [NCI*NCO][ctas_x][OUT_PER_CTA]
  float* chan_part;   // [NCO][ctas_x][2][8]   (only written by ci-chunk 0)
  int B, H, W, HO, WO, pad, vec_ok;
  float eps;
};

template <int CI, int CO, int K, int PAD>
__global__ void __launch_bounds__(WG_THREADS, 2) nconv_wgrad_tiled(const WGradArgs a) {
  using C = WGradCfg<CI, CO, K>;
  constexpr int CCH = C::CCH, CO_T = C::CO_T, RH = C::RH, P = C::P, NG = C::NG;
  extern __shared__ __align__(16) float smem[];
  float* s_gn = smem;                    // [CO_T][WG_TH][WG_TW]
  float* s_gd = s_gn + C::G_FLOATS;
  float* s_xc = s_gd + C::G_FLOATS;      // [CCH][RH][SW]
  float* s_c = s_xc + C::X_FLOATS;
  float* s_red = s_c + C::X_FLOATS;      // [P][CO_T][K]
  float* s_part = s_red + C::RED_FLOATS; // [8 warps][2][8]

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int chunk = blockIdx.y;                 // = oc * NCI + ic
  const int ic = chunk % C::NCI, oc = chunk / C::NCI;
  const int pair = tid % P, grp = tid / P;      // grp >= NG: idle in the FMA phase
  const int pi = pair / K, pu = pair % K;       // input channel (in chunk) and kernel row

  for (int i = tid; i < C::PART_FLOATS; i += WG_THREADS) s_part[i] = 0.f;

  // With an even CO_T the FMA stream is packed (FFMA2): gN / gD of a channel PAIR sit interleaved in shared
  // memory, so one 64-bit operand x one broadcast activation updates the taps of two output channels.
  constexpr bool PACK = (CO_T % 2 == 0);
  constexpr int CO_P = PACK ? CO_T / 2 : 1;
  float acc[PACK ? 1 : CO_T][K];
  u64 acc2[CO_P][K];
#pragma unroll
  for (int o = 0; o < (PACK ? 1 : CO_T); ++o)
#pragma unroll
    for (int v = 0; v < K; ++v) acc[o][v] = 0.f;
#pragma unroll
  for (int o = 0; o < CO_P; ++o)
#pragma unroll
    for (int v = 0; v < K; ++v) acc2[o][v] = 0ull;

  const int tiles_x = ceil_div(a.WO, WG_TW), tiles_y = ceil_div(a.HO, WG_TH);
  const int ntiles = a.B * tiles_y * tiles_x;
  constexpr int OFF = (4 - (PAD & 3)) & 3;   // staged col of the first needed element

#pragma unroll 1
  for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int txi = tile % tiles_x, tyi = (tile / tiles_x) % tiles_y, b = tile / (tiles_x * tiles_y);
    const int ox0 = txi * WG_TW, oy0 = tyi * WG_TH;
    __syncthreads();
    // ---- stage gN, gD for CO_T channels (and the per-channel sums for gb and the gcout term)
    // one slot = 4 pixels of one channel (scalar path) or of a channel PAIR (packed path: the pair of a pixel
    // is written as one 64-bit word, two conflict-free 128-bit stores per slot)
    auto load_channel = [&](int oo, int gy_, int gx_, float* n4, float* d4, float& sgy, float& sT) {
#pragma unroll
      for (int e = 0; e < 4; ++e) { n4[e] = 0.f; d4[e] = 0.f; }
      sgy = 0.f; sT = 0.f;
      if (gy_ < a.HO) {
        size_t base = ((size_t)(b * CO + oo) * a.HO + gy_) * a.WO;
        float g1[4] = {0.f, 0.f, 0.f, 0.f}, g2[4] = {0.f, 0.f, 0.f, 0.f}, yv[4] = {0.f, 0.f, 0.f, 0.f},
              cv[4] = {0.f, 0.f, 0.f, 0.f};
        bool in[4];
        if (a.vec_ok && gx_ + 3 < a.WO) {
          ld4(g1, a.gy + base + gx_); ld4(yv, a.y + base + gx_); ld4(cv, a.co + base + gx_);
          if (a.gco) ld4(g2, a.gco + base + gx_);
          in[0] = in[1] = in[2] = in[3] = true;
        } else {
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            in[e] = gx_ + e < a.WO;
            if (in[e]) {
              g1[e] = __ldg(a.gy + base + gx_ + e); yv[e] = __ldg(a.y + base + gx_ + e);
              cv[e] = __ldg(a.co + base + gx_ + e);
              if (a.gco) g2[e] = __ldg(a.gco + base + gx_ + e);
            }
          }
        }
        const float sw = __ldg(a.wsum + oo), bo = __ldg(a.bias + oo);
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          if (in[e]) {
            float inv = 1.f / (cv[e] * sw + a.eps);
            float gn = g1[e] * inv;
            n4[e] = gn;
            d4[e] = g2[e] / sw - gn * (yv[e] - bo);
            sgy += g1[e];
            sT += g2[e] * cv[e];
          }
        }
      }
    };
    constexpr int SLOT_CH = PACK ? 2 : 1;
    for (int idx = tid; idx < (CO_T / SLOT_CH) * WG_TH * (WG_TW / 4); idx += WG_THREADS) {
      int j = idx % (WG_TW / 4);
      int r = (idx / (WG_TW / 4)) % WG_TH;
      int os = idx / ((WG_TW / 4) * WG_TH);    // warp-uniform: 128 slots per channel (pair)
      const int gy_ = oy0 + r, gx_ = ox0 + 4 * j;
      float n4[SLOT_CH][4], d4[SLOT_CH][4], sgy[SLOT_CH], sT[SLOT_CH];
#pragma unroll
      for (int h = 0; h < SLOT_CH; ++h) load_channel(oc * CO_T + os * SLOT_CH + h, gy_, gx_, n4[h], d4[h], sgy[h], sT[h]);
      if constexpr (PACK) {
        float* pn = s_gn + ((os * WG_TH + r) * WG_TW + 4 * j) * 2;
        float* pd = s_gd + ((os * WG_TH + r) * WG_TW + 4 * j) * 2;
        *reinterpret_cast<float4*>(pn) = make_float4(n4[0][0], n4[1][0], n4[0][1], n4[1][1]);
        *reinterpret_cast<float4*>(pn + 4) = make_float4(n4[0][2], n4[1][2], n4[0][3], n4[1][3]);
        *reinterpret_cast<float4*>(pd) = make_float4(d4[0][0], d4[1][0], d4[0][1], d4[1][1]);
        *reinterpret_cast<float4*>(pd + 4) = make_float4(d4[0][2], d4[1][2], d4[0][3], d4[1][3]);
      } else {
        int so = (os * WG_TH + r) * WG_TW + 4 * j;
        *reinterpret_cast<float4*>(s_gn + so) = make_float4(n4[0][0], n4[0][1], n4[0][2], n4[0][3]);
        *reinterpret_cast<float4*>(s_gd + so) = make_float4(d4[0][0], d4[0][1], d4[0][2], d4[0][3]);
      }
      if (ic == 0) {
#pragma unroll
        for (int h = 0; h < SLOT_CH; ++h) {
          const float a1 = warp_sum(sgy[h]), a2 = warp_sum(sT[h]);
          const int o = os * SLOT_CH + h;
          if (lane == 0) { s_part[(warp * 2 + 0) * 8 + o] += a1; s_part[(warp * 2 + 1) * 8 + o] += a2; }
        }
      }
    }
    // ---- stage x*c and c halo tiles for CCH input channels
    const int gx0a = ox0 - PAD - OFF, gy0 = oy0 - PAD;
    if (a.vec_ok) {
      // 128-bit rows (W % 4 == 0, staged columns start at a multiple of 4: a quad is entirely inside or outside the
      // row): the loads of WS_UN slots are issued before the first is consumed, as in nconv_tiled's forward staging
      constexpr int WSLOTS = CCH * RH * (SW / 4), WNIT = (WSLOTS + WG_THREADS - 1) / WG_THREADS, WS_UN = 3;
#pragma unroll
      for (int it0 = 0; it0 < WNIT; it0 += WS_UN) {
        float4 xq[WS_UN], cq[WS_UN];
        int so[WS_UN];
#pragma unroll
        for (int k = 0; k < WS_UN; ++k) {
          const int idx = tid + (it0 + k) * WG_THREADS;
          const int j = idx % (SW / 4), r = (idx / (SW / 4)) % RH, cc = idx / ((SW / 4) * RH);
          const int gy_ = gy0 + r, gx_ = gx0a + 4 * j;
          so[k] = (it0 + k < WNIT && idx < WSLOTS) ? 4 * idx : -1;            // slots are in memory order
          const bool ok = so[k] >= 0 && gy_ >= 0 && gy_ < a.H && gx_ >= 0 && gx_ + 3 < a.W;
          xq[k] = make_float4(0.f, 0.f, 0.f, 0.f); cq[k] = xq[k];
          if (ok) {
            const size_t base = ((size_t)(b * CI + ic * CCH + cc) * a.H + gy_) * a.W + gx_;
            xq[k] = __ldg(reinterpret_cast<const float4*>(a.x + base));
            cq[k] = __ldg(reinterpret_cast<const float4*>(a.c + base));
          }
        }
#pragma unroll
        for (int k = 0; k < WS_UN; ++k) {
          if (so[k] < 0) continue;
          *reinterpret_cast<float4*>(s_xc + so[k]) = make_float4(xq[k].x * cq[k].x, xq[k].y * cq[k].y, xq[k].z * cq[k].z, xq[k].w * cq[k].w);
          *reinterpret_cast<float4*>(s_c + so[k]) = cq[k];
        }
      }
    } else
    for (int idx = tid; idx < CCH * RH * (SW / 4); idx += WG_THREADS) {
      int j = idx % (SW / 4);
      int r = (idx / (SW / 4)) % RH;
      int cc = idx / ((SW / 4) * RH);
      int gy_ = gy0 + r, gx_ = gx0a + 4 * j;
      float xv[4] = {0.f, 0.f, 0.f, 0.f}, cv[4] = {0.f, 0.f, 0.f, 0.f};
      if (gy_ >= 0 && gy_ < a.H) {
        size_t base = ((size_t)(b * CI + ic * CCH + cc) * a.H + gy_) * a.W;
        if (a.vec_ok && gx_ >= 0 && gx_ + 3 < a.W) { ld4(xv, a.x + base + gx_); ld4(cv, a.c + base + gx_); }
        else {
#pragma unroll
          for (int e = 0; e < 4; ++e)
            if (gx_ + e >= 0 && gx_ + e < a.W) { xv[e] = __ldg(a.x + base + gx_ + e); cv[e] = __ldg(a.c + base + gx_ + e); }
        }
      }
      int so = (cc * RH + r) * SW + 4 * j;
      *reinterpret_cast<float4*>(s_xc + so) = make_float4(xv[0] * cv[0], xv[1] * cv[1], xv[2] * cv[2], xv[3] * cv[3]);
      *reinterpret_cast<float4*>(s_c + so) = make_float4(cv[0], cv[1], cv[2], cv[3]);
    }
    __syncthreads();
    // ---- accumulate: thread (pi,pu) x all CO_T x K taps, over its strips of 4 pixels
    if (grp < NG) {
#pragma unroll 1
      for (int s = grp; s < WG_TH * (WG_TW / 4); s += NG) {
        const int r = s / (WG_TW / 4), j = s % (WG_TW / 4);
        float xr[12], cr[12];
        const float* px = s_xc + (pi * RH + r + pu) * SW + 4 * j;
        const float* pc = s_c + (pi * RH + r + pu) * SW + 4 * j;
#pragma unroll
        for (int q = 0; q < 3; ++q) { VecLoad<4>::ld(xr + 4 * q, px + 4 * q); VecLoad<4>::ld(cr + 4 * q, pc + 4 * q); }
        if constexpr (PACK) {
#pragma unroll
          for (int o = 0; o < CO_P; ++o) {
            const ulonglong2* pn = reinterpret_cast<const ulonglong2*>(s_gn + ((o * WG_TH + r) * WG_TW + 4 * j) * 2);
            const ulonglong2* pd = reinterpret_cast<const ulonglong2*>(s_gd + ((o * WG_TH + r) * WG_TW + 4 * j) * 2);
            const ulonglong2 n01 = pn[0], n23 = pn[1], d01 = pd[0], d23 = pd[1];
            const u64 gn2[4] = {n01.x, n01.y, n23.x, n23.y}, gd2[4] = {d01.x, d01.y, d23.x, d23.y};
#pragma unroll
            for (int v = 0; v < K; ++v)
#pragma unroll
              for (int p = 0; p < 4; ++p) {
                ffma2(acc2[o][v], gn2[p], xr[OFF + p + v]);
                ffma2(acc2[o][v], gd2[p], cr[OFF + p + v]);
              }
          }
        } else {
#pragma unroll
          for (int o = 0; o < CO_T; ++o) {
            float gn[4], gd[4];
            VecLoad<4>::ld(gn, s_gn + (o * WG_TH + r) * WG_TW + 4 * j);
            VecLoad<4>::ld(gd, s_gd + (o * WG_TH + r) * WG_TW + 4 * j);
#pragma unroll
            for (int v = 0; v < K; ++v)
#pragma unroll
              for (int p = 0; p < 4; ++p) {
                acc[o][v] = fmaf(gn[p], xr[OFF + p + v], acc[o][v]);
                acc[o][v] = fmaf(gd[p], cr[OFF + p + v], acc[o][v]);
              }
          }
        }
      }
    }
  }

  // ---- fixed-order reduction over the strip groups, then one partial per CTA
  for (int i = tid; i < C::RED_FLOATS; i += WG_THREADS) s_red[i] = 0.f;
  __syncthreads();
  for (int g = 0; g < NG; ++g) {
    if (grp == g) {
#pragma unroll
      for (int o = 0; o < CO_T; ++o)
#pragma unroll
        for (int v = 0; v < K; ++v) {
          const float av = PACK ? ((o & 1) ? hi32f(acc2[PACK ? o >> 1 : 0][v]) : lo32f(acc2[PACK ? o >> 1 : 0][v])) : acc[PACK ? 0 : o][v];
          s_red[(pair * CO_T + o) * K + v] += av;
        }
    }
    __syncthreads();
  }
  float* dst = a.partial + ((size_t)chunk * gridDim.x + blockIdx.x) * C::OUT_PER_CTA;
  for (int i = tid; i < C::OUT_PER_CTA; i += WG_THREADS) {
    // i -> (o, cc, u, v) in weight order; s_red is [(cc*K+u)][o][v]
    int v = i % K, u = (i / K) % K, cc = (i / (K * K)) % CCH, o = i / (K * K * CCH);
    dst[i] = s_red[((cc * K + u) * CO_T + o) * K + v];
  }
  if (ic == 0 && tid < 16) {
    int which = tid / 8, o = tid % 8;
    float t = 0.f;
    for (int w8 = 0; w8 < 8; ++w8) t += s_part[(w8 * 2 + which) * 8 + o];
    a.chan_part[(((size_t)oc * gridDim.x + blockIdx.x) * 2 + which) * 8 + o] = t;
  }
}

template <int CI, int CO, int K>
__global__ void nconv_wgrad_final(const float* __restrict__ partial, const float* __restrict__ chan_part,
                                  const float* __restrict__ wsum, float* __restrict__ gw,
                                  float* __restrict__ gb, int ctas_x) {
  // one warp per output element; lane l sums partials l, l+32, ... in a fixed order, then a
  // fixed-shape shuffle tree: deterministic for a given ctas_x
  using C = WGradCfg<CI, CO, K>;
  const int total = CO * CI * K * K;
  const int lane = threadIdx.x & 31;
  const int wid = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = (gridDim.x * blockDim.x) >> 5;
  for (int idx = wid; idx < total + CO; idx += nw) {
    if (idx < total) {
      int v = idx % K, u = (idx / K) % K, i = (idx / (K * K)) % CI, o = idx / (K * K * CI);
      int oc = o / C::CO_T, ol = o % C::CO_T, ic = i / C::CCH, il = i % C::CCH;
      int chunk = oc * C::NCI + ic;
      int local = ((ol * C::CCH + il) * K + u) * K + v;
      float s = 0.f, T = 0.f;
      for (int cta = lane; cta < ctas_x; cta += 32) {
        s += __ldg(partial + ((size_t)chunk * ctas_x + cta) * C::OUT_PER_CTA + local);
        T += __ldg(chan_part + (((size_t)oc * ctas_x + cta) * 2 + 1) * 8 + ol);
      }
      s = warp_sum(s); T = warp_sum(T);
      if (lane == 0) gw[idx] = s - T / __ldg(wsum + o);
    } else if (gb) {
      int o = idx - total;
      int oc = o / C::CO_T, ol = o % C::CO_T;
      float s = 0.f;
      for (int cta = lane; cta < ctas_x; cta += 32) s += __ldg(chan_part + (((size_t)oc * ctas_x + cta) * 2 + 0) * 8 + ol);
      s = warp_sum(s);
      if (lane == 0) gb[o] = s;
    }
  }
}

// ---- generic fallbacks (any shape; simple, not tuned) ------------------------------------------
__global__ void nconv_bwd_data_generic(const float* __restrict__ gy, const float* __restrict__ gco,
                                       const float* __restrict__ x, const float* __restrict__ c,
                                       const float* __restrict__ w, const float* __restrict__ bias,
                                       const float* __restrict__ y, const float* __restrict__ co,
                                       const float* __restrict__ wsum, float* __restrict__ gx,
                                       float* __restrict__ gc, int B, int Ci, int Co, int H, int W,
                                       int kh, int kw, int ph, int pw, int HO, int WO, float eps) {
  long long total = (long long)B * Ci * H * W;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    int qx = (int)(idx % W);
    long long r = idx / W;
    int qy = (int)(r % H); r /= H;
    int i = (int)(r % Ci);
    int b = (int)(r / Ci);
    float a1 = 0.f, a2 = 0.f;
    for (int o = 0; o < Co; ++o) {
      const float sw = __ldg(wsum + o), bo = __ldg(bias + o);
      for (int u = 0; u < kh; ++u) {
        int py = qy + ph - u;
        if (py < 0 || py >= HO) continue;
        for (int v = 0; v < kw; ++v) {
          int px = qx + pw - v;
          if (px < 0 || px >= WO) continue;
          size_t off = (((size_t)b * Co + o) * HO + py) * WO + px;
          float inv = 1.f / (__ldg(co + off) * sw + eps);
          float gn = __ldg(gy + off) * inv;
          float gd = (gco ? __ldg(gco + off) : 0.f) / sw - gn * (__ldg(y + off) - bo);
          float wv = __ldg(w + ((o * Ci + i) * kh + u) * kw + v);
          a1 = fmaf(wv, gn, a1);
          a2 = fmaf(wv, gd, a2);
        }
      }
    }
    if (gx) gx[idx] = a1 * __ldg(c + idx);
    if (gc) gc[idx] = fmaf(a1, __ldg(x + idx), a2);
  }
}

// one block per weight element (plus Co blocks for the bias gradient)
__global__ void nconv_bwd_weight_generic(const float* __restrict__ gy, const float* __restrict__ gco,
                                         const float* __restrict__ x, const float* __restrict__ c,
                                         const float* __restrict__ bias, const float* __restrict__ y,
                                         const float* __restrict__ co, const float* __restrict__ wsum,
                                         float* __restrict__ gw, float* __restrict__ gb, int B, int Ci,
                                         int Co, int H, int W, int kh, int kw, int ph, int pw, int HO,
                                         int WO, float eps) {
  __shared__ float red[2][8];
  const int nw = Co * Ci * kh * kw;
  const int idx = blockIdx.x;
  const bool is_w = idx < nw;
  int o, i = 0, u = 0, v = 0;
  if (is_w) { v = idx % kw; u = (idx / kw) % kh; i = (idx / (kw * kh)) % Ci; o = idx / (kw * kh * Ci); }
  else o = idx - nw;
  const float sw = __ldg(wsum + o), bo = __ldg(bias + o);
  float s = 0.f, T = 0.f;
  long long total = (long long)B * HO * WO;
  for (long long t = threadIdx.x; t < total; t += blockDim.x) {
    int px = (int)(t % WO);
    int py = (int)((t / WO) % HO);
    int b = (int)(t / ((long long)WO * HO));
    size_t off = (((size_t)b * Co + o) * HO + py) * WO + px;
    float g1 = __ldg(gy + off);
    if (!is_w) { s += g1; continue; }
    float g2 = gco ? __ldg(gco + off) : 0.f;
    float cv = __ldg(co + off);
    T += g2 * cv;
    int qy = py - ph + u, qx = px - pw + v;
    if (qy < 0 || qy >= H || qx < 0 || qx >= W) continue;
    float inv = 1.f / (cv * sw + eps);
    float gn = g1 * inv;
    float gd = g2 / sw - gn * (__ldg(y + off) - bo);
    size_t ioff = (((size_t)b * Ci + i) * H + qy) * W + qx;
    float ci = __ldg(c + ioff);
    s = fmaf(gn, __ldg(x + ioff) * ci, s);
    s = fmaf(gd, ci, s);
  }
  s = warp_sum(s); T = warp_sum(T);
  if ((threadIdx.x & 31) == 0) { red[0][threadIdx.x >> 5] = s; red[1][threadIdx.x >> 5] = T; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float ss = 0.f, TT = 0.f;
    for (int k = 0; k < (int)(blockDim.x >> 5); ++k) { ss += red[0][k]; TT += red[1][k]; }
    if (is_w) gw[idx] = ss - TT / sw;
    else if (gb) gb[o] = ss;
  }
}

static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }
constexpr int WG_CTAS_X_MAX = 148 * 6;

template <int CI, int CO, int K, int PAD>
static int launch_wgrad(WGradArgs a, float* ws_partial, float* gw, float* gb, cudaStream_t st) {
  using C = WGradCfg<CI, CO, K>;
  auto kern = nconv_wgrad_tiled<CI, CO, K, PAD>;
  static bool attr_done[NCONV_MAX_DEVICES] = {};
  int dev = 0;
  NCONV_RETURN_IF_ERR(cudaGetDevice(&dev));
  if (C::BYTES > 48 * 1024 && (dev >= NCONV_MAX_DEVICES || !attr_done[dev])) {
    NCONV_RETURN_IF_ERR(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::BYTES));
    if (dev < NCONV_MAX_DEVICES) attr_done[dev] = true;
  }
  const int chunks = C::NCI * C::NCO;
  const int ntiles = a.B * ceil_div(a.HO, WG_TH) * ceil_div(a.WO, WG_TW);
  static int occ_cache[NCONV_MAX_DEVICES] = {};
  int occ = dev < NCONV_MAX_DEVICES ? occ_cache[dev] : 0;
  if (occ == 0) {
    NCONV_RETURN_IF_ERR(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, WG_THREADS, C::BYTES));
    if (occ < 1) occ = 1;
    if (occ > 6) occ = 6;
    if (dev < NCONV_MAX_DEVICES) occ_cache[dev] = occ;
  }
  int ctas_x = (occ * 148 + chunks - 1) / chunks;
  if (ctas_x > WG_CTAS_X_MAX) ctas_x = WG_CTAS_X_MAX;
  if (ctas_x > ntiles) ctas_x = ntiles;
  a.partial = ws_partial;
  a.chan_part = ws_partial + (size_t)chunks * ctas_x * C::OUT_PER_CTA;
  char label[64];
  snprintf(label, sizeof(label), "nconv_wgrad %d->%d k%d", CI, CO, K);
  {
    // reads gy, gcout, y, cout once per input-channel chunk and x, c once per output chunk
    const double pin = (double)a.B * a.H * a.W, pout = (double)a.B * a.HO * a.WO;
    LaunchScope ls(st, label, 4.0 * CI * CO * K * K * pout, 4.0 * (2.0 * CI * pin + 4.0 * CO * pout));
    kern<<<dim3(ctas_x, chunks), WG_THREADS, C::BYTES, st>>>(a);
  }
  NCONV_RETURN_IF_ERR(cudaGetLastError());
  const int total = CO * CI * K * K + CO;
  {
    LaunchScope ls(st, "nconv_wgrad_final", 0.0, 4.0 * (double)chunks * ctas_x * C::OUT_PER_CTA);
    nconv_wgrad_final<CI, CO, K><<<ceil_div(total, 4), 128, 0, st>>>(a.partial, a.chan_part, a.wsum, gw, gb, ctas_x);
  }
  return (int)cudaGetLastError();
}

// partial-buffer floats needed by the tiled weight gradient (upper bound over instantiations)
static size_t wgrad_partial_floats(int Ci, int Co, int K) {
  int cch = Ci < 4 ? Ci : 4, cot = Co < 8 ? Co : 8;
  int chunks = (Ci / cch) * (Co / cot);
  size_t ctas = WG_CTAS_X_MAX;
  return (size_t)chunks * ctas * ((size_t)cot * cch * K * K) + (size_t)(Co / cot) * ctas * 16;
}

}  // namespace nconv

using namespace nconv;

extern "C" size_t nconv_b200_nconv2d_bwd_workspace_bytes(int B, int Ci, int Co, int H, int W, int kh,
                                                         int kw, int ph, int pw) {
  (void)B; (void)H; (void)W; (void)ph; (void)pw;
  if (Ci <= 0 || Co <= 0 || kh <= 0 || kw <= 0) return 0;
  size_t fl = align_up((size_t)Co, 64);
  int K = kh > kw ? kh : kw;
  fl += wgrad_partial_floats(Ci, Co, K) + 64;
  if (Co == 2 && Ci <= 2 && K == 3) { size_t bs = (size_t)BS_MAX_UNITS * (Co * Ci * 9 + 2 * Co) + 64; if (bs > fl) fl = bs; }
  if (Co == 2 && Ci <= 2 && K == 5) { size_t bs = (size_t)BS_MAX_UNITS * (Co * 25 + 2 * Co) + 64; if (bs > fl) fl = bs; }
  return fl * sizeof(float);
}

#define BWD_DATA_CASE(cof_, cif_, cot_, k_, padf_)                                                   \
  if (Co == cof_ && Ci == cif_ && kh == k_ && ph == padf_)                                           \
    return launch_tiled<cof_, 0, cif_, cot_, k_, (k_ - 1 - padf_), F_BWD>(t, st);
#define WGRAD_CASE(cif_, cof_, k_, padf_)                                                            \
  if (Ci == cif_ && Co == cof_ && kh == k_ && wa.pad == padf_)                                       \
    return launch_wgrad<cif_, cof_, k_, padf_>(wa, part, gw, gb, st);

static int bwd_data_dispatch(const TiledArgs& t, int Ci, int Co, int kh, int ph, cudaStream_t st) {
  BWD_DATA_CASE(8, 8, 8, 5, 2)
  BWD_DATA_CASE(8, 16, 8, 3, 1)
  BWD_DATA_CASE(8, 16, 8, 3, 0)
  BWD_DATA_CASE(1, 8, 8, 1, 2)
  BWD_DATA_CASE(8, 1, 1, 5, 2)
  BWD_DATA_CASE(2, 1, 1, 3, 1)
  BWD_DATA_CASE(2, 2, 2, 3, 1)
  BWD_DATA_CASE(4, 4, 4, 3, 1)
  BWD_DATA_CASE(16, 16, 8, 3, 1)
  BWD_DATA_CASE(2, 1, 1, 5, 2)
  BWD_DATA_CASE(2, 2, 2, 5, 2)
  BWD_DATA_CASE(4, 4, 4, 5, 2)
  BWD_DATA_CASE(16, 16, 8, 5, 2)
  return NCONV_ERR_UNSUPPORTED;
}

static int wgrad_dispatch(const WGradArgs& wa, float* part, float* gw, float* gb, int Ci, int Co,
                          int kh, cudaStream_t st) {
  WGRAD_CASE(1, 8, 5, 2)
  WGRAD_CASE(8, 8, 5, 2)
  WGRAD_CASE(16, 8, 3, 1)
  WGRAD_CASE(16, 8, 3, 0)
  WGRAD_CASE(8, 1, 1, 2)
  WGRAD_CASE(1, 2, 3, 1)
  WGRAD_CASE(2, 2, 3, 1)
  WGRAD_CASE(4, 4, 3, 1)
  WGRAD_CASE(16, 16, 3, 1)
  WGRAD_CASE(1, 2, 5, 2)
  WGRAD_CASE(2, 2, 5, 2)
  WGRAD_CASE(4, 4, 5, 2)
  WGRAD_CASE(16, 16, 5, 2)
  return NCONV_ERR_UNSUPPORTED;
}

extern "C" int nconv_b200_nconv2d_bwd(const float* gy, const float* gcout, const float* x,
                                      const float* c, const float* w, const float* bias,
                                      const float* y, const float* cout, float* gx, float* gc,
                                      float* gw, float* gb, int B, int Ci, int Co, int H, int W,
                                      int kh, int kw, int ph, int pw, float eps, void* workspace,
                                      size_t workspace_bytes, void* stream) {
  if (!gy || !x || !c || !w || !bias || !y || !cout || !workspace) return NCONV_ERR_BAD_ARG;
  if (B <= 0 || Ci <= 0 || Co <= 0 || H <= 0 || W <= 0 || kh <= 0 || kw <= 0 || ph < 0 || pw < 0)
    return NCONV_ERR_BAD_ARG;
  const int HO = H + 2 * ph - kh + 1, WO = W + 2 * pw - kw + 1;
  if (HO <= 0 || WO <= 0) return NCONV_ERR_BAD_ARG;
  if (workspace_bytes < nconv_b200_nconv2d_bwd_workspace_bytes(B, Ci, Co, H, W, kh, kw, ph, pw))
    return NCONV_ERR_WORKSPACE;
  cudaStream_t st = (cudaStream_t)stream;
  float* wsum = (float*)workspace;
  float* part = wsum + align_up((size_t)Co, 64);

  // HBM-bound small-channel shapes (BASELINE.json config 2): one fused streaming pass (nconv_bwd_stream.cuh).
  // NCONV_B200_NO_BWD_STREAM=1 keeps them on the tiled kernels (A/B measurements only).
  {
    static const bool no_stream = getenv("NCONV_B200_NO_BWD_STREAM") != nullptr;
    const bool vec = (W % 4 == 0) && aligned16(gy) && aligned16(y) && aligned16(cout) && (!gcout || aligned16(gcout)) &&
                     aligned16(x) && aligned16(c) && (!gx || aligned16(gx)) && (!gc || aligned16(gc));
    if (!no_stream && gx && gc && gw && vec && kh == 3 && kw == 3 && ph == 1 && pw == 1 && Co == 2 && (Ci == 1 || Ci == 2)) {
      BsArgs a = {};
      a.gy = gy; a.gco = gcout; a.y = y; a.co = cout; a.x = x; a.c = c; a.w = w; a.bias = bias;
      a.gx = gx; a.gc = gc; a.partial = (float*)workspace; a.B = B; a.H = H; a.W = W; a.eps = eps;
      const int rc = (Ci == 1) ? launch_bwd_stream<1, 2>(a, gw, gb, st) : launch_bwd_stream<2, 2>(a, gw, gb, st);
      if (rc != NCONV_ERR_UNSUPPORTED) return rc;
    }
    if (!no_stream && gx && gc && gw && vec && kh == 5 && kw == 5 && ph == 2 && pw == 2 && Co == 2 && (Ci == 1 || Ci == 2)) {
      BsArgs a = {};
      a.gy = gy; a.gco = gcout; a.y = y; a.co = cout; a.x = x; a.c = c; a.w = w; a.bias = bias;
      a.gx = gx; a.gc = gc; a.partial = (float*)workspace; a.B = B; a.H = H; a.W = W; a.eps = eps;
      const int rc = (Ci == 1) ? launch_bwd_stream5<1>(a, gw, gb, st) : launch_bwd_stream5<2>(a, gw, gb, st);
      if (rc != NCONV_ERR_UNSUPPORTED) return rc;
    }
  }

  {
    LaunchScope ls(st, "wsum", 0.0, 4.0 * Co * Ci * kh * kw);
    wsum_kernel<<<Co, 32, 0, st>>>(w, wsum, Co, Ci * kh * kw);
  }
  NCONV_RETURN_IF_ERR(cudaGetLastError());

  const bool square = (kh == kw) && (ph == pw);
  const bool vec_ok = (W % 4 == 0) && (WO % 4 == 0) && aligned16(gy) && aligned16(y) && aligned16(cout) &&
                      (!gcout || aligned16(gcout)) && aligned16(x) && aligned16(c) &&
                      (!gx || aligned16(gx)) && (!gc || aligned16(gc));
  if (gx || gc) {
    int rc = NCONV_ERR_UNSUPPORTED;
    if (square) {
      TiledArgs t = {};
      t.xa = gy; t.ca = gcout; t.xb = y; t.cb = cout; t.w = w; t.bias = bias;
      t.y = gx; t.co = gc; t.ex = x; t.ec = c; t.wsum = wsum;
      t.B = B; t.H = HO; t.W = WO; t.HO = H; t.WO = W; t.a_first = 1; t.vec_ok = vec_ok; t.eps = eps;
      t.map_h = NearestMap::make(HO, HO); t.map_w = NearestMap::make(WO, WO);
      rc = bwd_data_dispatch(t, Ci, Co, kh, ph, st);
    }
    if (rc == NCONV_ERR_UNSUPPORTED) {
      long long total = (long long)B * Ci * H * W;
      int blocks = (int)((total + 255) / 256);
      if (blocks > 148 * 32) blocks = 148 * 32;
      {
        LaunchScope ls(st, "nconv_bwd_data_generic", 4.0 * Ci * Co * kh * kw * (double)B * H * W,
                       4.0 * B * (4.0 * Ci * H * W + 4.0 * Co * HO * WO));
        nconv_bwd_data_generic<<<blocks, 256, 0, st>>>(gy, gcout, x, c, w, bias, y, cout, wsum, gx, gc, B,
                                                       Ci, Co, H, W, kh, kw, ph, pw, HO, WO, eps);
      }
      rc = (int)cudaGetLastError();
    }
    if (rc) return rc;
  }
  if (gw) {
    int rc = NCONV_ERR_UNSUPPORTED;
    if (square) {
      WGradArgs wa = {};
      wa.gy = gy; wa.gco = gcout; wa.y = y; wa.co = cout; wa.x = x; wa.c = c; wa.bias = bias;
      wa.wsum = wsum; wa.B = B; wa.H = H; wa.W = W; wa.HO = HO; wa.WO = WO; wa.pad = ph;
      wa.vec_ok = vec_ok; wa.eps = eps;
      rc = wgrad_dispatch(wa, part, gw, gb, Ci, Co, kh, st);
    }
    if (rc == NCONV_ERR_UNSUPPORTED) {
      {
        LaunchScope ls(st, "nconv_bwd_weight_generic", 4.0 * Ci * Co * kh * kw * (double)B * HO * WO,
                       4.0 * B * (2.0 * Ci * H * W + 4.0 * Co * HO * WO));
        nconv_bwd_weight_generic<<<Co * Ci * kh * kw + Co, 256, 0, st>>>(gy, gcout, x, c, bias, y, cout, wsum,
                                                                         gw, gb, B, Ci, Co, H, W, kh, kw,
                                                                         ph, pw, HO, WO, eps);
      }
      rc = (int)cudaGetLastError();
    }
    if (rc) return rc;
  }
  return 0;
}
