// Dual-stream tiled correlation core shared by the NConv2d forward and backward-data kernels.
//
// Forward (/root/reference/models/step1.py:116-149): streams (x*c, c) correlated with W give
// (N, D); epilogue y = N/(D+eps)+b, cout = D/sum(W).  Flags fuse the neighbouring DNET stages
// (step1.py:51-94): `S > 0.01` threshold, the two independent 2x2 max-pools, nearest-upsample +
// concat on the load side, and the nconv7 1x1 + crop tail.
//
// Backward-data: streams (gN, gD) — formed at load time from (gy, gcout, y,
// cout) — correlated with the flipped/transposed W give (g_xc, g_cD); epilogue
// gx = g_xc*c, gc = g_xc*x + g_cD.
//
// Design (SIMT; these layers have 1..16 channels, see DESIGN.md for why tensor cores do not pay
// here): a CTA of 256 threads owns a 16x64 output tile.  Staged channels go through shared
// memory in chunks of <=4 as halo tiles of the two streams, filled with 128-bit coalesced loads;
// zero padding is produced by the fill.  Each thread accumulates a 1x4 strip of pixels for CO_T
// output channels of BOTH correlations in registers (2*4*CO_T accumulators); weights are
// transposed into shared memory once per CTA as [ci][kh][kw][co] so the inner loop reads them as
// broadcast 128-bit LDS and the FFMA stream is a register outer product.
#pragma once
#include "common.cuh"
#include "nconv_b200.h"
#include "launch_count.h"
#include <cstdio>

namespace nconv {

constexpr int TW = 64;        // output tile width
constexpr int TH = 16;        // output tile height
constexpr int PX = 4;         // pixels per thread (one row strip)
constexpr int NTHREADS = 256; // (TW/PX) x TH
constexpr int SW = TW + 8;    // staged row width (floats): tile + aligned halo

constexpr int F_BWD = 64;     // internal flag: backward-data mode

struct TiledArgs {
  // forward: source A (xa, ca) [B,CA,H,W]; source B (xb, cb) [B,CB,HB,WB] nearest-upsampled
  // backward: xa = gy, ca = gcout, xb = y, cb = cout, all [B,CA,H,W] (forward-output space)
  const float* xa; const float* ca; const float* xb; const float* cb;
  const float* w; const float* bias;
  // forward: y, co [B,CO,HO,WO] (+ pooled yp, cp).  backward: y = gx, co = gc [B,CO,HO,WO]
  float* y; float* co; float* yp; float* cp;
  const float* w7; const float* b7; float* out;   // tail
  const float* ex; const float* ec;               // backward epilogue inputs x, c [B,CO,HO,WO]
  const float* wsum;                              // backward: sum(W[o]) per forward out channel
  int B, H, W, HB, WB, HO, WO, HP, WP;
  int a_first;   // weight input-channel order: A first (1) or B first (0)
  int vec_ok;    // 128-bit global access allowed (W%4==0, WO%4==0, aligned bases)
  int up2_vec;   // source B is an exact 2x upsample with 8-byte-aligned rows (W == 2*WB, WB even, aligned bases)
  float eps, thr;
  NearestMap map_h, map_w;
};

template <int N> struct VecLoad;
template <> struct VecLoad<4> {
  __device__ static __forceinline__ void ld(float* d, const float* s) {
    float4 v = *reinterpret_cast<const float4*>(s); d[0] = v.x; d[1] = v.y; d[2] = v.z; d[3] = v.w; }
};
template <> struct VecLoad<2> {
  __device__ static __forceinline__ void ld(float* d, const float* s) {
    float2 v = *reinterpret_cast<const float2*>(s); d[0] = v.x; d[1] = v.y; }
};
template <> struct VecLoad<1> {
  __device__ static __forceinline__ void ld(float* d, const float* s) { d[0] = s[0]; }
};

template <int CA, int CB, int CO_T, int K>
struct TiledSmem {
  static constexpr int CI = CA + CB;
  static constexpr int CCH = CI < 4 ? CI : 4;
  static constexpr int RH = TH + K - 1;
  static constexpr int W_FLOATS = ((CI * K * K * CO_T) + 3) & ~3;
  static constexpr int S_FLOATS = 8;  // per-co weight sums / bias (CO_T <= 8)
  static constexpr int TILE_FLOATS = CCH * RH * SW;
  static constexpr size_t BYTES = sizeof(float) * (W_FLOATS + 2 * S_FLOATS + 2 * TILE_FLOATS);
};

__device__ __forceinline__ void ld4(float* d, const float* p) {
  float4 t = __ldg(reinterpret_cast<const float4*>(p));
  d[0] = t.x; d[1] = t.y; d[2] = t.z; d[3] = t.w;
}

// CA/CB: staged (reduction) channels from source A / B.  CO: channels produced, CO_T per CTA.
// PAD: zero padding of the correlation (may be negative in backward of an over-padded layer).
template <int CA, int CB, int CO, int CO_T, int K, int PAD, int FLAGS>
__global__ void __launch_bounds__(NTHREADS, (CO_T * PX * 2 <= 64) ? 2 : 1)
nconv_tiled(const TiledArgs a) {
  using SM = TiledSmem<CA, CB, CO_T, K>;
  constexpr int CI = SM::CI, CCH = SM::CCH, RH = SM::RH;
  constexpr bool THRESH = (FLAGS & NCONV_B200_F_THRESH) != 0;
  constexpr bool POOL = (FLAGS & NCONV_B200_F_POOL) != 0;
  constexpr bool TAIL = (FLAGS & NCONV_B200_F_TAIL) != 0;
  constexpr bool BWD = (FLAGS & F_BWD) != 0;
  constexpr int OFF = (4 - (PAD & 3)) & 3;           // staged col of the first needed element
  constexpr int NV = (OFF + PX + K - 1 + 3) / 4;     // float4s per thread per staged row
  constexpr int WV = (CO_T % 4 == 0) ? 4 : (CO_T % 2 == 0 ? 2 : 1);
  static_assert(CO % CO_T == 0 && CO_T <= 8, "bad CO tiling");
  static_assert(!TAIL || (CO_T == CO), "tail needs all channels in one thread");
  static_assert(!BWD || CB == 0, "backward has a single source");
  static_assert(4 * NV + 4 * (TW / PX - 1) <= SW, "staged row too narrow");

  extern __shared__ __align__(16) float smem[];
  float* s_w = smem;                       // [CI][K][K][CO_T]
  float* s_sum = s_w + SM::W_FLOATS;       // [CO_T] 1 / sum of weights per output channel
  float* s_bias = s_sum + SM::S_FLOATS;    // [CO_T]
  float* s_xc = s_bias + SM::S_FLOATS;     // [CCH][RH][SW]  stream 1 (x*c | gN)
  float* s_c = s_xc + SM::TILE_FLOATS;     // [CCH][RH][SW]  stream 2 (c   | gD)

  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  constexpr int NG = CO / CO_T;
  const int b = blockIdx.z / NG, cog = blockIdx.z % NG;
  const int ox0 = blockIdx.x * TW, oy0 = blockIdx.y * TH;

  // ---- prologue: weights -> smem as [staged channel][u][v][o].
  for (int idx = tid; idx < CI * K * K * CO_T; idx += NTHREADS) {
    int o = idx % CO_T;
    int r = idx / CO_T;
    int v = r % K; r /= K;
    int u = r % K; r /= K;
    int cs = r;
    float wv;
    if (BWD) {
      // staged channel = forward output channel, o = forward input channel, taps flipped
      wv = __ldg(&a.w[(((size_t)cs * CO + cog * CO_T + o) * K + (K - 1 - u)) * K + (K - 1 - v)]);
    } else {
      // staged order is A then B whatever the concat order of the reference layer was
      int ciw;
      if (CB == 0) ciw = cs;
      else if (a.a_first) ciw = cs;
      else ciw = (cs < CA) ? (CB + cs) : (cs - CA);
      wv = __ldg(&a.w[(((size_t)(cog * CO_T + o) * CI + ciw) * K + u) * K + v]);
    }
    s_w[idx] = wv;
  }
  __syncthreads();
  if (!BWD) {
    const int warp = tid >> 5, lane = tid & 31;
    for (int o = warp; o < CO_T; o += NTHREADS / 32) {
      float s = 0.f;
      for (int j = lane; j < CI * K * K; j += 32) s += s_w[j * CO_T + o];
      s = warp_sum(s);
      if (lane == 0) { s_sum[o] = 1.f / s; s_bias[o] = __ldg(&a.bias[cog * CO_T + o]); }   // reciprocal: cout = D * (1/sum W)
    }
  }

  // Accumulators.  With an even CO_T the FMA stream is packed: one FFMA2 (fma.rn.f32x2, sm_100)
  // updates two adjacent output channels of one pixel — the weight pair (w[o], w[o+1]) comes
  // straight from a 64-bit half of the broadcast LDS.128 and the activation is the scalar
  // broadcast operand — so the 2*PX*CO_T FMAs per tap cost PX*CO_T issue slots.  fp32 peak needs
  // every issue slot with scalar FFMA (measured 72 TFLOP/s), which a loop that also issues LDS and
  // address arithmetic cannot give; FFMA2 runs at the same flop rate on half the slots.
  constexpr bool PACK = (CO_T % 2 == 0);
  constexpr int CO_P = PACK ? CO_T / 2 : 1;
  float accN[PACK ? 1 : CO_T][PX], accD[PACK ? 1 : CO_T][PX];
  u64 accN2[CO_P][PX], accD2[CO_P][PX];
  if constexpr (PACK) {
#pragma unroll
    for (int o = 0; o < CO_P; ++o)
#pragma unroll
      for (int p = 0; p < PX; ++p) { accN2[o][p] = 0ull; accD2[o][p] = 0ull; }
  } else {
#pragma unroll
    for (int o = 0; o < CO_T; ++o)
#pragma unroll
      for (int p = 0; p < PX; ++p) { accN[o][p] = 0.f; accD[o][p] = 0.f; }
  }

  const int gx0a = ox0 - PAD - OFF;   // global col of staged col 0 (multiple of 4)
  const int gy0 = oy0 - PAD;

#pragma unroll 1
  for (int c0 = 0; c0 < CI; c0 += CCH) {
    __syncthreads();
    // ---- stage CCH channels: halo tiles of both streams, zero-filled outside the image
    constexpr int SLOTS = CCH * RH * (SW / 4);
    // Fast staging (forward, 128-bit path, chunk entirely from one source): the loads of STAGE_UN slots are issued
    // before any of them is consumed, so a chunk costs ~NIT/STAGE_UN DRAM round trips instead of NIT (the generic
    // loop below is load -> multiply -> STS per slot and its branches keep the compiler from hoisting the loads).
    // With W % 4 == 0 and staged columns starting at a multiple of 4, a quad is entirely inside or outside the row.
    bool staged = false;
    if constexpr (!BWD) {
      const bool from_a = (CB == 0) || (c0 + CCH <= CA);
      const bool from_b = (CB > 0) && (c0 >= CA);
      if (a.vec_ok && (from_a || (from_b && a.up2_vec))) {
        staged = true;
        constexpr int NIT = (SLOTS + NTHREADS - 1) / NTHREADS;
        constexpr int STAGE_UN = 3;
#pragma unroll
        for (int it0 = 0; it0 < NIT; it0 += STAGE_UN) {
          float4 xv[STAGE_UN], cv[STAGE_UN];
          int so[STAGE_UN];
          bool ok[STAGE_UN];
#pragma unroll
          for (int k = 0; k < STAGE_UN; ++k) {
            const int idx = tid + (it0 + k) * NTHREADS;
            const int j = idx % (SW / 4), r = (idx / (SW / 4)) % RH, cc = idx / ((SW / 4) * RH);
            const int cs = c0 + cc, gy = gy0 + r, gx = gx0a + 4 * j;
            so[k] = (it0 + k < NIT && idx < SLOTS) ? (cc * RH + r) * SW + 4 * j : -1;
            ok[k] = so[k] >= 0 && gy >= 0 && gy < a.H && gx >= 0 && gx + 3 < a.W && cs < CI;
            xv[k] = make_float4(0.f, 0.f, 0.f, 0.f); cv[k] = xv[k];
            if (ok[k]) {
              if (from_a) {
                const size_t base = ((size_t)(b * CA + cs) * a.H + gy) * a.W + gx;
                xv[k] = __ldg(reinterpret_cast<const float4*>(a.xa + base));
                if (!THRESH) cv[k] = __ldg(reinterpret_cast<const float4*>(a.ca + base));
              } else {
                const size_t base = ((size_t)(b * CB + (cs - CA)) * a.HB + a.map_h(gy)) * a.WB + (gx >> 1);
                const float2 x2 = __ldg(reinterpret_cast<const float2*>(a.xb + base));
                const float2 c2 = __ldg(reinterpret_cast<const float2*>(a.cb + base));
                xv[k] = make_float4(x2.x, x2.x, x2.y, x2.y);
                cv[k] = make_float4(c2.x, c2.x, c2.y, c2.y);
              }
            }
          }
#pragma unroll
          for (int k = 0; k < STAGE_UN; ++k) {
            if (so[k] < 0) continue;
            float4 c4 = cv[k];
            if (THRESH && from_a) {
              c4.x = (ok[k] && xv[k].x > a.thr) ? 1.f : 0.f; c4.y = (ok[k] && xv[k].y > a.thr) ? 1.f : 0.f;
              c4.z = (ok[k] && xv[k].z > a.thr) ? 1.f : 0.f; c4.w = (ok[k] && xv[k].w > a.thr) ? 1.f : 0.f;
            }
            *reinterpret_cast<float4*>(s_xc + so[k]) = make_float4(xv[k].x * c4.x, xv[k].y * c4.y, xv[k].z * c4.z, xv[k].w * c4.w);
            *reinterpret_cast<float4*>(s_c + so[k]) = c4;
          }
        }
      }
    }
    if (!staged)
    for (int idx = tid; idx < SLOTS; idx += NTHREADS) {
      int j = idx % (SW / 4);
      int r = (idx / (SW / 4)) % RH;
      int cc = idx / ((SW / 4) * RH);
      int cs = c0 + cc;
      int gy = gy0 + r, gx = gx0a + 4 * j;
      float s1[4] = {0.f, 0.f, 0.f, 0.f}, s2[4] = {0.f, 0.f, 0.f, 0.f};
      if (gy >= 0 && gy < a.H && cs < CI) {
        if (BWD) {
          // gN = gy/(D+eps); gD = -gy*(y-b)/(D+eps) + gcout/s, with D = cout*s
          size_t base = ((size_t)(b * CA + cs) * a.H + gy) * a.W;
          float g1[4] = {0.f, 0.f, 0.f, 0.f}, g2[4] = {0.f, 0.f, 0.f, 0.f}, yv[4] = {0.f, 0.f, 0.f, 0.f},
                cv[4] = {0.f, 0.f, 0.f, 0.f};
          bool in[4];
          if (a.vec_ok && gx >= 0 && gx + 3 < a.W) {
            ld4(g1, a.xa + base + gx); ld4(yv, a.xb + base + gx); ld4(cv, a.cb + base + gx);
            if (a.ca) ld4(g2, a.ca + base + gx);
            in[0] = in[1] = in[2] = in[3] = true;
          } else {
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              int g = gx + e;
              in[e] = (g >= 0 && g < a.W);
              if (in[e]) {
                g1[e] = __ldg(a.xa + base + g); yv[e] = __ldg(a.xb + base + g);
                cv[e] = __ldg(a.cb + base + g);
                if (a.ca) g2[e] = __ldg(a.ca + base + g);
              }
            }
          }
          const float sw = __ldg(a.wsum + cs), bo = __ldg(a.bias + cs);
          const float rsw = 1.f / sw;
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            if (in[e]) {
              float gn = __fdividef(g1[e], cv[e] * sw + a.eps);
              s1[e] = gn;
              s2[e] = g2[e] * rsw - gn * (yv[e] - bo);
            }
          }
        } else if (cs < CA) {
          size_t base = ((size_t)(b * CA + cs) * a.H + gy) * a.W;
          float xv[4] = {0.f, 0.f, 0.f, 0.f};
          if (a.vec_ok && gx >= 0 && gx + 3 < a.W) {
            ld4(xv, a.xa + base + gx);
            if (!THRESH) ld4(s2, a.ca + base + gx);
          } else {
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              int g = gx + e;
              if (g >= 0 && g < a.W) {
                xv[e] = __ldg(a.xa + base + g);
                if (!THRESH) s2[e] = __ldg(a.ca + base + g);
              }
            }
          }
          if (THRESH) {
#pragma unroll
            for (int e = 0; e < 4; ++e) s2[e] = (xv[e] > a.thr) ? 1.f : 0.f;
          }
#pragma unroll
          for (int e = 0; e < 4; ++e) s1[e] = xv[e] * s2[e];
        } else if (CB > 0) {
          int sy = a.map_h(gy);
          size_t base = ((size_t)(b * CB + (cs - CA)) * a.HB + sy) * a.WB;
          if (a.up2_vec && gx >= 0 && gx + 3 < a.W) {
            // exact 2x nearest upsample: 4 output px = 2 source px = one 8-byte load per plane
            const float2 x2 = __ldg(reinterpret_cast<const float2*>(a.xb + base + (gx >> 1)));
            const float2 c2 = __ldg(reinterpret_cast<const float2*>(a.cb + base + (gx >> 1)));
            s2[0] = c2.x; s2[1] = c2.x; s2[2] = c2.y; s2[3] = c2.y;
            s1[0] = x2.x * c2.x; s1[1] = s1[0]; s1[2] = x2.y * c2.y; s1[3] = s1[2];
          } else
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            int g = gx + e;
            if (g >= 0 && g < a.W) {
              int sx = a.map_w(g);
              float xv = __ldg(a.xb + base + sx);
              s2[e] = __ldg(a.cb + base + sx);
              s1[e] = xv * s2[e];
            }
          }
        }
      }
      int so = (cc * RH + r) * SW + 4 * j;
      *reinterpret_cast<float4*>(s_xc + so) = make_float4(s1[0], s1[1], s1[2], s1[3]);
      *reinterpret_cast<float4*>(s_c + so) = make_float4(s2[0], s2[1], s2[2], s2[3]);
    }
    __syncthreads();

    // ---- accumulate both correlations for this chunk
    const int nch = (CI - c0) < CCH ? (CI - c0) : CCH;
#pragma unroll 1
    for (int cc = 0; cc < nch; ++cc) {
      const float* wrow = s_w + (size_t)(c0 + cc) * K * K * CO_T;
#pragma unroll
      for (int u = 0; u < K; ++u) {
        float xr[4 * NV], cr[4 * NV];
        const int so = (cc * RH + ty + u) * SW + 4 * tx;
#pragma unroll
        for (int q = 0; q < NV; ++q) {
          VecLoad<4>::ld(xr + 4 * q, s_xc + so + 4 * q);
          VecLoad<4>::ld(cr + 4 * q, s_c + so + 4 * q);
        }
#pragma unroll
        for (int v = 0; v < K; ++v) {
          if constexpr (PACK) {
            u64 w2[CO_P];
            if constexpr (CO_T % 4 == 0) {
#pragma unroll
              for (int o = 0; o < CO_P; o += 2) {
                const ulonglong2 t = *reinterpret_cast<const ulonglong2*>(wrow + (u * K + v) * CO_T + 2 * o);
                w2[o] = t.x; w2[o + 1] = t.y;
              }
            } else {
#pragma unroll
              for (int o = 0; o < CO_P; ++o) w2[o] = *reinterpret_cast<const u64*>(wrow + (u * K + v) * CO_T + 2 * o);
            }
#pragma unroll
            for (int o = 0; o < CO_P; ++o)
#pragma unroll
              for (int p = 0; p < PX; ++p) {
                ffma2(accN2[o][p], w2[o], xr[OFF + p + v]);
                ffma2(accD2[o][p], w2[o], cr[OFF + p + v]);
              }
          } else {
            float wv[CO_T];
#pragma unroll
            for (int o = 0; o < CO_T; o += WV) VecLoad<WV>::ld(wv + o, wrow + (u * K + v) * CO_T + o);
#pragma unroll
            for (int o = 0; o < CO_T; ++o)
#pragma unroll
              for (int p = 0; p < PX; ++p) {
                accN[o][p] = fmaf(wv[o], xr[OFF + p + v], accN[o][p]);
                accD[o][p] = fmaf(wv[o], cr[OFF + p + v], accD[o][p]);
              }
          }
        }
      }
    }
  }

  // ---- epilogue
  float aN[CO_T][PX], aD[CO_T][PX];
#pragma unroll
  for (int o = 0; o < CO_T; ++o)
#pragma unroll
    for (int p = 0; p < PX; ++p) {
      if constexpr (PACK) {
        aN[o][p] = (o & 1) ? hi32f(accN2[o >> 1][p]) : lo32f(accN2[o >> 1][p]);
        aD[o][p] = (o & 1) ? hi32f(accD2[o >> 1][p]) : lo32f(accD2[o >> 1][p]);
      } else { aN[o][p] = accN[o][p]; aD[o][p] = accD[o][p]; }
    }
  const int oy = oy0 + ty, ox = ox0 + PX * tx;
  if (BWD) {
    if (oy >= a.HO) return;
#pragma unroll
    for (int o = 0; o < CO_T; ++o) {
      size_t base = (((size_t)b * CO + cog * CO_T + o) * a.HO + oy) * a.WO + ox;
      float xv[4] = {0.f, 0.f, 0.f, 0.f}, cv[4] = {0.f, 0.f, 0.f, 0.f};
      const bool vec = a.vec_ok && ox + 3 < a.WO;
      if (vec) { ld4(xv, a.ex + base); ld4(cv, a.ec + base); }
      else {
#pragma unroll
        for (int p = 0; p < PX; ++p)
          if (ox + p < a.WO) { xv[p] = __ldg(a.ex + base + p); cv[p] = __ldg(a.ec + base + p); }
      }
      float gxv[PX], gcv[PX];
#pragma unroll
      for (int p = 0; p < PX; ++p) {
        gxv[p] = aN[o][p] * cv[p];
        gcv[p] = fmaf(aN[o][p], xv[p], aD[o][p]);
      }
      if (vec) {
        if (a.y) *reinterpret_cast<float4*>(a.y + base) = make_float4(gxv[0], gxv[1], gxv[2], gxv[3]);
        if (a.co) *reinterpret_cast<float4*>(a.co + base) = make_float4(gcv[0], gcv[1], gcv[2], gcv[3]);
      } else {
#pragma unroll
        for (int p = 0; p < PX; ++p)
          if (ox + p < a.WO) {
            if (a.y) a.y[base + p] = gxv[p];
            if (a.co) a.co[base + p] = gcv[p];
          }
      }
    }
    return;
  }
  if (TAIL) {
    // nconv6 (valid 3x3, evaluated centred) -> nconv7 (1x1 over CO channels, its pad-2 ring
    // of zeros gives exactly bias7 on the border) -> crop.  step1.py:90-94.
    if (oy < a.H) {
      const float b7 = __ldg(a.b7);
      float res[PX];
#pragma unroll
      for (int p = 0; p < PX; ++p) {
        float n7 = 0.f, d7 = 0.f;
#pragma unroll
        for (int o = 0; o < CO_T; ++o) {
          float D = aD[o][p];
          float x6 = __fdividef(aN[o][p], D + a.eps) + s_bias[o];
          float c6 = D * s_sum[o];
          float w7 = __ldg(a.w7 + o);
          n7 = fmaf(w7, x6 * c6, n7);
          d7 = fmaf(w7, c6, d7);
        }
        int gx = ox + p;
        bool ring = (oy == 0) || (oy == a.H - 1) || (gx == 0) || (gx >= a.W - 1);
        res[p] = ring ? b7 : (__fdividef(n7, d7 + a.eps) + b7);
      }
      size_t base = ((size_t)b * a.H + oy) * a.W + ox;
      if (a.vec_ok && ox + 3 < a.W) {
        *reinterpret_cast<float4*>(a.out + base) = make_float4(res[0], res[1], res[2], res[3]);
      } else {
#pragma unroll
        for (int p = 0; p < PX; ++p) if (ox + p < a.W) a.out[base + p] = res[p];
      }
    }
    return;
  }

  float yv[CO_T][PX], cv[CO_T][PX];
#pragma unroll
  for (int o = 0; o < CO_T; ++o) {
    const float bo = s_bias[o], so = s_sum[o];
#pragma unroll
    for (int p = 0; p < PX; ++p) {
      float D = aD[o][p];
      // fast division (MUFU.RCP + FMUL, 2 ulp): the IEEE sequence cost as much as 10 % (8->8 k5) to 45 %
      // (1->8 k5) of the FMA loop; the parity bound is 1e-4
      yv[o][p] = __fdividef(aN[o][p], D + a.eps) + bo;
      cv[o][p] = D * so;
    }
  }
  if (oy < a.HO) {
#pragma unroll
    for (int o = 0; o < CO_T; ++o) {
      size_t base = (((size_t)b * CO + cog * CO_T + o) * a.HO + oy) * a.WO + ox;
      if (a.vec_ok && ox + 3 < a.WO) {
        *reinterpret_cast<float4*>(a.y + base) = make_float4(yv[o][0], yv[o][1], yv[o][2], yv[o][3]);
        *reinterpret_cast<float4*>(a.co + base) = make_float4(cv[o][0], cv[o][1], cv[o][2], cv[o][3]);
      } else {
#pragma unroll
        for (int p = 0; p < PX; ++p)
          if (ox + p < a.WO) { a.y[base + p] = yv[o][p]; a.co[base + p] = cv[o][p]; }
      }
    }
  }
  if (POOL) {
    // Independent 2x2/2 max-pools of y and of cout (step1.py:62-74).  Window order is
    // row-major with ATen's `(v > m) || isnan(v)` update, so the value (and NaN behaviour)
    // is exactly F.max_pool2d's.  The lower row of each window lives in lane^16.
    const int py = oy >> 1;
#pragma unroll
    for (int o = 0; o < CO_T; ++o) {
      float ylo[PX], clo[PX];
#pragma unroll
      for (int p = 0; p < PX; ++p) {
        ylo[p] = __shfl_xor_sync(0xffffffffu, yv[o][p], 16);
        clo[p] = __shfl_xor_sync(0xffffffffu, cv[o][p], 16);
      }
      if ((ty & 1) == 0 && py < a.HP) {
        size_t base = (((size_t)b * CO + cog * CO_T + o) * a.HP + py) * a.WP;
#pragma unroll
        for (int h = 0; h < PX / 2; ++h) {
          int pxo = (ox >> 1) + h;
          if (pxo < a.WP) {
            float m = -INFINITY;
            m = pool_max(m, yv[o][2 * h]); m = pool_max(m, yv[o][2 * h + 1]);
            m = pool_max(m, ylo[2 * h]);   m = pool_max(m, ylo[2 * h + 1]);
            float n = -INFINITY;
            n = pool_max(n, cv[o][2 * h]); n = pool_max(n, cv[o][2 * h + 1]);
            n = pool_max(n, clo[2 * h]);   n = pool_max(n, clo[2 * h + 1]);
            a.yp[base + pxo] = m;
            a.cp[base + pxo] = n;
          }
        }
      }
    }
  }
}

template <int CA, int CB, int CO, int CO_T, int K, int PAD, int FLAGS>
static int launch_tiled(const TiledArgs& a, cudaStream_t st) {
  using SM = TiledSmem<CA, CB, CO_T, K>;
  auto kern = nconv_tiled<CA, CB, CO, CO_T, K, PAD, FLAGS>;
  static bool attr_done[NCONV_MAX_DEVICES] = {};
  if (SM::BYTES > 48 * 1024) {
    int dev = 0;
    NCONV_RETURN_IF_ERR(cudaGetDevice(&dev));
    if (dev >= NCONV_MAX_DEVICES || !attr_done[dev]) {
      NCONV_RETURN_IF_ERR(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                               (int)SM::BYTES));
      if (dev < NCONV_MAX_DEVICES) attr_done[dev] = true;
    }
  }
  const bool tail = (FLAGS & NCONV_B200_F_TAIL) != 0;
  int ow = tail ? a.W : a.WO, oh = tail ? a.H : a.HO;
  dim3 grid(ceil_div(ow, TW), ceil_div(oh, TH), a.B * (CO / CO_T));
  // Algorithmic work of this launch.  Forward: both streams of both sources are read once
  // (the low-resolution source at its own size, one plane instead of two under F_THRESH) and
  // every output written once.  Backward-data: gy, gcout, y, cout in; x, c in; gx, gc out.
  constexpr bool BWD = (FLAGS & F_BWD) != 0;
  constexpr bool THRESH = (FLAGS & NCONV_B200_F_THRESH) != 0;
  constexpr bool POOL = (FLAGS & NCONV_B200_F_POOL) != 0;
  const double px_in = (double)a.B * a.H * a.W, px_out = (double)a.B * oh * ow;
  double bytes, flops;
  if (BWD) {
    bytes = 4.0 * (4.0 * CA * px_in + 4.0 * CO * px_out);
    flops = 4.0 * CA * CO * K * K * px_out;
  } else if (tail) {
    bytes = 4.0 * (2.0 * CA * px_in + 2.0 * CB * a.B * a.HB * a.WB + px_out);
    flops = 4.0 * (CA + CB) * CO * K * K * px_out + 4.0 * CO * px_out;
  } else {
    bytes = 4.0 * ((THRESH ? 1.0 : 2.0) * CA * px_in + 2.0 * CB * a.B * a.HB * a.WB + 2.0 * CO * px_out +
                   (POOL ? 2.0 * CO * a.B * a.HP * a.WP : 0.0));
    flops = 4.0 * (CA + CB) * CO * K * K * px_out;
  }
  char label[64];
  snprintf(label, sizeof(label), "%s %d+%d->%d k%d f%d", BWD ? "nconv_bwd_data" : "nconv_fwd", CA, CB, CO, K,
           FLAGS & 7);
  {
    LaunchScope ls(st, label, flops, bytes);
    kern<<<grid, NTHREADS, SM::BYTES, st>>>(a);
  }
  return (int)cudaGetLastError();
}

}  // namespace nconv
