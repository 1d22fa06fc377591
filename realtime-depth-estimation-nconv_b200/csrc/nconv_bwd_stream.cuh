// Streaming NConv2d backward for the HBM-bound small-channel shapes (Ci in {1,2}, Co = 2, K = 3, pad 1):
// data gradient, weight gradient and bias gradient in ONE pass over the tensors.
// Math: nconv_bwd.cu header (derived from /root/reference/models/step1.py:116-149, which the reference
// differentiates with autograd).  Same schedule idea as nconv_stream.cuh:
//   * one WARP owns a 128-column strip x RS rows of one image and walks it top to bottom; no shared-memory
//     staging, no block barrier in the loop; rows are prefetched one step ahead into registers;
//   * per output-space row the lane turns (gy, gcout, y, cout) into the packed pair (gN, gD) of its 4 pixels,
//     gets the one-column halo from the neighbouring lanes by shuffle (strip-edge lanes load theirs) and keeps
//     the windows of the last three rows in registers;
//   * input row q then needs exactly those three windows, for BOTH results:
//       A[i][q]            = sum_{o,u,v} W[o,i,u,v] * (gN, gD)[o][q + 1 - (u,v)]      (FFMA2 on the pair)
//       gx = A.N * c,  gc = A.N * x + A.D
//       gW[o,i,u,v]       += (x*c)[i][q] * gN[o][q + 1 - (u,v)] + c[i][q] * gD[o][q + 1 - (u,v)]
//     so every byte of the six tensors is read once and gx, gc are written once:
//     algorithmic bytes 4*B*(4*Co + 4*Ci)*H*W.
//   * per-unit partial sums of gW, gb and T[o] = sum gcout*cout go to a workspace; a second tiny kernel adds
//     them in a fixed order (deterministic) and applies  gW -= T[o] / sum(W[o]).
#pragma once
#include "common.cuh"
#include "nconv_b200.h"
#include "launch_count.h"
#include "nconv_stream.cuh"

namespace nconv {

constexpr int BS_WARPS = 8;
constexpr int BS_MAX_UNITS = 8192;     // workspace rows reserved for per-unit partials

struct BsArgs {
  const float* gy; const float* gco; const float* y; const float* co;
  const float* x; const float* c; const float* w; const float* bias;
  float* gx; float* gc;
  float* partial;            // [units][NP]
  int B, H, W, RS, nstrip, nseg;
  long long units;
  float eps;
};

// acc.{lo,hi} += a.{lo,hi} * b.{lo,hi}
__device__ __forceinline__ void ffma2pp(u64& acc, u64 a, u64 b) {
  asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(acc) : "l"(a), "l"(b));
}
__device__ __forceinline__ u64 pack2f(float lo, float hi) {
  return (u64)__float_as_uint(lo) | ((u64)__float_as_uint(hi) << 32);
}
__device__ __forceinline__ u64 shfl_up64(u64 v) {
  const unsigned lo = __shfl_up_sync(0xffffffffu, (unsigned)v, 1), hi = __shfl_up_sync(0xffffffffu, (unsigned)(v >> 32), 1);
  return (u64)lo | ((u64)hi << 32);
}
__device__ __forceinline__ u64 shfl_down64(u64 v) {
  const unsigned lo = __shfl_down_sync(0xffffffffu, (unsigned)v, 1), hi = __shfl_down_sync(0xffffffffu, (unsigned)(v >> 32), 1);
  return (u64)lo | ((u64)hi << 32);
}

template <int CO>
struct BsRawG {               // one prefetched output-space row of a lane
  float4 gy[CO], gco[CO], y[CO], co[CO];
  float hgy[CO], hgco[CO], hy[CO], hco[CO];      // strip-edge halo column (lanes 0 and 31 only)
};
template <int CI>
struct BsRawX { float4 x[CI], c[CI]; };

template <int CO, bool HAS_GCO>
__device__ __forceinline__ void bs_load_g(BsRawG<CO>& r, const BsArgs& a, long long off, size_t plane, int hoff, bool own, bool halo) {
  const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
  for (int o = 0; o < CO; ++o) {
    r.gy[o] = z; r.gco[o] = z; r.y[o] = z; r.co[o] = z;
    r.hgy[o] = 0.f; r.hgco[o] = 0.f; r.hy[o] = 0.f; r.hco[o] = 0.f;
    if (own) {
      r.gy[o] = __ldg(reinterpret_cast<const float4*>(a.gy + off + o * plane));
      if (HAS_GCO) r.gco[o] = __ldg(reinterpret_cast<const float4*>(a.gco + off + o * plane));
      r.y[o] = __ldg(reinterpret_cast<const float4*>(a.y + off + o * plane));
      r.co[o] = __ldg(reinterpret_cast<const float4*>(a.co + off + o * plane));
    }
    if (halo) {
      r.hgy[o] = __ldg(a.gy + off + o * plane + hoff);
      if (HAS_GCO) r.hgco[o] = __ldg(a.gco + off + o * plane + hoff);
      r.hy[o] = __ldg(a.y + off + o * plane + hoff);
      r.hco[o] = __ldg(a.co + off + o * plane + hoff);
    }
  }
}

template <int CI>
__device__ __forceinline__ void bs_load_x(BsRawX<CI>& r, const BsArgs& a, long long off, size_t plane, bool own) {
  const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
  for (int i = 0; i < CI; ++i) {
    r.x[i] = z; r.c[i] = z;
    if (own) {
      r.x[i] = __ldg(reinterpret_cast<const float4*>(a.x + off + i * plane));
      r.c[i] = __ldg(reinterpret_cast<const float4*>(a.c + off + i * plane));
    }
  }
}

template <int CI, int CO, bool HAS_GCO>
__global__ void __launch_bounds__(BS_WARPS * 32, 1) nconv_bwd_stream(const BsArgs a) {
  constexpr int NW = CO * CI * 9, NP = NW + 2 * CO;
  __shared__ float s_w[NW], s_is[CO], s_b[CO];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int i = tid; i < NW; i += BS_WARPS * 32) s_w[i] = __ldg(a.w + i);
  __syncthreads();
  if (tid < CO) {
    float s = 0.f;
    for (int j = 0; j < CI * 9; ++j) s += s_w[tid * CI * 9 + j];
    s_is[tid] = 1.f / s;
    s_b[tid] = __ldg(a.bias + tid);
  }
  __syncthreads();

  const long long unit = (long long)blockIdx.x * BS_WARPS + warp;
  if (unit >= a.units) return;
  const int strip = (int)(unit % a.nstrip);
  const int seg = (int)((unit / a.nstrip) % a.nseg);
  const int b = (int)(unit / ((long long)a.nseg * a.nstrip));
  const int r0 = seg * a.RS, x0 = strip * ST_COLS, col = x0 + ST_PX * lane;
  const int rows = min(a.RS, a.H - r0);
  const int W = a.W, H = a.H;
  const size_t plane = (size_t)H * W;
  const bool inb = col < W;
  const int hcol = (lane == 0) ? x0 - 1 : x0 + ST_COLS;
  const bool hlane = (lane == 0 || lane == 31) && hcol >= 0 && hcol < W;
  const int hoff = (lane == 0) ? -1 : ST_PX;
  float sden[CO], isum[CO], bo[CO];
#pragma unroll
  for (int o = 0; o < CO; ++o) { isum[o] = s_is[o]; sden[o] = 1.f / isum[o]; bo[o] = s_b[o]; }

  // element offsets of (channel 0, row, own column) in the output-space / input-space tensors
  long long goff = ((long long)b * CO * H + (r0 - 1)) * W + col;      // next g row to load
  long long xoff = ((long long)b * CI * H + r0) * W + col;            // next x row to load
  long long ooff = xoff;                                              // next gx/gc row to store

  u64 win[3][CO][6];                    // (gN, gD) windows of the last three output-space rows
  float accW[CO][CI][9], accB[CO], accT[CO];
#pragma unroll
  for (int o = 0; o < CO; ++o) {
    accB[o] = 0.f; accT[o] = 0.f;
#pragma unroll
    for (int i = 0; i < CI; ++i)
#pragma unroll
      for (int t = 0; t < 9; ++t) accW[o][i][t] = 0.f;
  }

  BsRawG<CO> rg;
  BsRawX<CI> rx;
  {
    const bool ok = r0 - 1 >= 0;
    bs_load_g<CO, HAS_GCO>(rg, a, goff, plane, hoff, ok && inb, ok && hlane);
    goff += W;
    bs_load_x<CI>(rx, a, xoff, plane, inb);        // row r0: consumed at step 2
    xoff += W;
  }

  const int nsteps = rows + 2;
#pragma unroll 1
  for (int sb = 0; sb < nsteps; sb += 3) {
#pragma unroll
    for (int sl = 0; sl < 3; ++sl) {
      const int s = sb + sl;
      if (s < nsteps) {
        // ---- output-space row gr = r0-1+s: (gy, gcout, y, cout) -> packed (gN, gD), window with halo
        {
          const int gr = r0 - 1 + s;
          const bool count = (s >= 1) && (s <= rows);            // rows of the segment proper: gb and T sums
#pragma unroll
          for (int o = 0; o < CO; ++o) {
            const float gy4[4] = {rg.gy[o].x, rg.gy[o].y, rg.gy[o].z, rg.gy[o].w};
            const float gc4[4] = {rg.gco[o].x, rg.gco[o].y, rg.gco[o].z, rg.gco[o].w};
            const float y4[4] = {rg.y[o].x, rg.y[o].y, rg.y[o].z, rg.y[o].w};
            const float c4[4] = {rg.co[o].x, rg.co[o].y, rg.co[o].z, rg.co[o].w};
#pragma unroll
            for (int p = 0; p < 4; ++p) {
              const float gn = gy4[p] * rcp_fast(fmaf(c4[p], sden[o], a.eps));
              const float gd = fmaf(gc4[p], isum[o], -gn * (y4[p] - bo[o]));
              win[sl][o][p + 1] = pack2f(gn, gd);
              if (count) { accB[o] += gy4[p]; accT[o] = fmaf(gc4[p], c4[p], accT[o]); }
            }
            const float hgn = rg.hgy[o] * rcp_fast(fmaf(rg.hco[o], sden[o], a.eps));
            const float hgd = fmaf(rg.hgco[o], isum[o], -hgn * (rg.hy[o] - bo[o]));
            const u64 hp = pack2f(hgn, hgd);
            const u64 l = shfl_up64(win[sl][o][4]), r = shfl_down64(win[sl][o][1]);
            win[sl][o][0] = (lane == 0) ? hp : l;
            win[sl][o][5] = (lane == 31) ? hp : r;
          }
          // prefetch the next output-space row
          const int gn_row = gr + 1;
          const bool ok = (s + 1 < nsteps) && gn_row >= 0 && gn_row < H;
          bs_load_g<CO, HAS_GCO>(rg, a, goff, plane, hoff, ok && inb, ok && hlane);
          goff += W;
        }
        // ---- input row q = r0+s-2: data gradient and weight-gradient accumulation
        if (s >= 2) {
          float xo[CI][4], cv[CI][4];
#pragma unroll
          for (int i = 0; i < CI; ++i) {
            xo[i][0] = rx.x[i].x; xo[i][1] = rx.x[i].y; xo[i][2] = rx.x[i].z; xo[i][3] = rx.x[i].w;
            cv[i][0] = rx.c[i].x; cv[i][1] = rx.c[i].y; cv[i][2] = rx.c[i].z; cv[i][3] = rx.c[i].w;
          }
          // prefetch the next input row (consumed at step s+1)
          bs_load_x<CI>(rx, a, xoff, plane, inb && (s + 1 < nsteps));
          xoff += W;
#pragma unroll
          for (int i = 0; i < CI; ++i) {
            u64 A[4] = {0ull, 0ull, 0ull, 0ull};
            float xc[4];
#pragma unroll
            for (int p = 0; p < 4; ++p) xc[p] = xo[i][p] * cv[i][p];
#pragma unroll
            for (int o = 0; o < CO; ++o)
#pragma unroll
              for (int u = 0; u < 3; ++u) {
                const int slot = (sl - u + 3) % 3;               // output-space row q+1-u
#pragma unroll
                for (int v = 0; v < 3; ++v) {
                  const float wv = s_w[((o * CI + i) * 3 + u) * 3 + v];
                  float acc = accW[o][i][u * 3 + v];
#pragma unroll
                  for (int p = 0; p < 4; ++p) {
                    const u64 g2 = win[slot][o][p + 2 - v];
                    ffma2(A[p], g2, wv);
                    acc = fmaf(lo32f(g2), xc[p], acc);
                    acc = fmaf(hi32f(g2), cv[i][p], acc);
                  }
                  accW[o][i][u * 3 + v] = acc;
                }
              }
            if (inb) {
              float gx4[4], gc4[4];
#pragma unroll
              for (int p = 0; p < 4; ++p) {
                gx4[p] = lo32f(A[p]) * cv[i][p];
                gc4[p] = fmaf(lo32f(A[p]), xo[i][p], hi32f(A[p]));
              }
              __stcs(reinterpret_cast<float4*>(a.gx + ooff + i * plane), make_float4(gx4[0], gx4[1], gx4[2], gx4[3]));
              __stcs(reinterpret_cast<float4*>(a.gc + ooff + i * plane), make_float4(gc4[0], gc4[1], gc4[2], gc4[3]));
            }
          }
          ooff += W;
        }
      }
    }
  }

  // ---- per-unit partial sums (fixed shuffle tree), one row of the workspace per unit
  float* dst = a.partial + (size_t)unit * NP;
#pragma unroll
  for (int o = 0; o < CO; ++o) {
#pragma unroll
    for (int i = 0; i < CI; ++i)
#pragma unroll
      for (int t = 0; t < 9; ++t) {
        const float v = warp_sum(accW[o][i][t]);
        if (lane == 0) dst[(o * CI + i) * 9 + t] = v;
      }
    const float vb = warp_sum(accB[o]), vt = warp_sum(accT[o]);
    if (lane == 0) { dst[NW + o] = vb; dst[NW + CO + o] = vt; }
  }
}

// gW[idx] = sum_units partial - T[o] / sum(W[o]);  gb[o] = sum_units partial.  One warp per output.
template <int CI, int CO>
__global__ void nconv_bwd_stream_final(const float* __restrict__ partial, const float* __restrict__ w,
                                       float* __restrict__ gw, float* __restrict__ gb, int units) {
  constexpr int NW = CO * CI * 9, NP = NW + 2 * CO;
  const int lane = threadIdx.x & 31;
  const int idx = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (idx >= NW + CO) return;
  if (idx < NW) {
    const int o = idx / (CI * 9);
    float s = 0.f, T = 0.f, ws = 0.f;
    for (int u = lane; u < units; u += 32) {
      s += __ldg(partial + (size_t)u * NP + idx);
      T += __ldg(partial + (size_t)u * NP + NW + CO + o);
    }
    for (int j = lane; j < CI * 9; j += 32) ws += __ldg(w + o * CI * 9 + j);
    s = warp_sum(s); T = warp_sum(T); ws = warp_sum(ws);
    if (lane == 0) gw[idx] = s - T / ws;
  } else if (gb) {
    const int o = idx - NW;
    float s = 0.f;
    for (int u = lane; u < units; u += 32) s += __ldg(partial + (size_t)u * NP + NW + o);
    s = warp_sum(s);
    if (lane == 0) gb[o] = s;
  }
}

template <int CI, int CO>
static int launch_bwd_stream(BsArgs a, float* gw, float* gb, cudaStream_t st) {
  constexpr int NW = CO * CI * 9;
  int dev = 0, sms = 0;
  NCONV_RETURN_IF_ERR(cudaGetDevice(&dev));
  static int sm_cache[NCONV_MAX_DEVICES] = {};
  if (dev < NCONV_MAX_DEVICES) sms = sm_cache[dev];
  if (sms == 0) {
    NCONV_RETURN_IF_ERR(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    if (dev < NCONV_MAX_DEVICES) sm_cache[dev] = sms;
  }
  // rows per unit: all units resident in whole waves (one CTA of 8 warps per SM); a unit walks RS + 2 rows
  const int nstrip = ceil_div(a.W, ST_COLS);
  const double cap = (double)sms * BS_WARPS;
  int best_rs = 16; double best = 1e300;
  for (int rs = 4; rs <= 128; ++rs) {
    const double units = (double)a.B * nstrip * ceil_div(a.H, rs);
    if (units > BS_MAX_UNITS) continue;
    const double cost = ceil(units / cap) * (rs + 2);
    if (cost < best - 1e-9) { best = cost; best_rs = rs; }
  }
  if (best >= 1e300) return NCONV_ERR_UNSUPPORTED;
  if (const char* e = getenv("NCONV_B200_BWD_STREAM_RS")) { int v = atoi(e); if (v > 0) best_rs = v; }
  a.RS = best_rs; a.nstrip = nstrip; a.nseg = ceil_div(a.H, best_rs);
  a.units = (long long)a.B * a.nstrip * a.nseg;
  if (a.units > BS_MAX_UNITS) return NCONV_ERR_UNSUPPORTED;
  const unsigned blocks = (unsigned)((a.units + BS_WARPS - 1) / BS_WARPS);
  const double px = (double)a.B * a.H * a.W;
  char label[64];
  snprintf(label, sizeof(label), "nconv_bwd_stream %d->%d k3", CI, CO);
  {
    LaunchScope ls(st, label, 8.0 * CI * CO * 9 * px, 4.0 * (4.0 * CO + 4.0 * CI) * px);
    if (a.gco) nconv_bwd_stream<CI, CO, true><<<blocks, BS_WARPS * 32, 0, st>>>(a);
    else nconv_bwd_stream<CI, CO, false><<<blocks, BS_WARPS * 32, 0, st>>>(a);
  }
  NCONV_RETURN_IF_ERR(cudaGetLastError());
  {
    LaunchScope ls(st, "nconv_bwd_stream_final", 0.0, 4.0 * (double)a.units * (NW + 2 * CO));
    nconv_bwd_stream_final<CI, CO><<<ceil_div((NW + CO) * 32, 128), 128, 0, st>>>(a.partial, a.w, gw, gb, (int)a.units);
  }
  return (int)cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------------------------
// The same single pass for K = 5 (pad 2), Co = 2, Ci in {1, 2} (BASELINE config 2's 1->2 / 2->2 k5 rows; before: two
// passes, nconv_bwd_data + nconv_wgrad, 0.23 / 0.27 ms at 16x480x640).  What changes against the 3x3 kernel:
//   * a lane owns 2 pixels (64-column strips): five window rows of 2 + 4 columns for both output channels are 120
//     registers, 4-pixel lanes would not fit;
//   * a unit handles ONE input channel (50 weight-gradient accumulators); with Ci = 2 the two units of a strip are
//     adjacent warps of one CTA, so the second one's (gy, gcout, y, cout) loads hit L1 / L2; bias and T sums are taken
//     by the channel-0 unit;
//   * two halo columns per side: from the neighbouring lane's two pixels by shuffle, strip-edge lanes load theirs.
// ---------------------------------------------------------------------------------------------------------------------
constexpr int B5_PX = 2, B5_COLS = 32 * B5_PX, B5_K = 5, B5_CO = 2, B5_WIN = B5_PX + 4;

struct B5RawG {
  float2 gy[B5_CO], gco[B5_CO], y[B5_CO], co[B5_CO];
  float2 hgy[B5_CO], hgco[B5_CO], hy[B5_CO], hco[B5_CO];      // strip-edge halo columns (lanes 0 and 31 only)
};

template <bool HAS_GCO>
__device__ __forceinline__ void b5_load_g(B5RawG& r, const BsArgs& a, long long off, size_t plane, int hoff, bool own, bool h0,
                                          bool h1) {
  const float2 z = make_float2(0.f, 0.f);
#pragma unroll
  for (int o = 0; o < B5_CO; ++o) {
    r.gy[o] = z; r.gco[o] = z; r.y[o] = z; r.co[o] = z;
    r.hgy[o] = z; r.hgco[o] = z; r.hy[o] = z; r.hco[o] = z;
    if (own) {
      r.gy[o] = __ldg(reinterpret_cast<const float2*>(a.gy + off + o * plane));
      if (HAS_GCO) r.gco[o] = __ldg(reinterpret_cast<const float2*>(a.gco + off + o * plane));
      r.y[o] = __ldg(reinterpret_cast<const float2*>(a.y + off + o * plane));
      r.co[o] = __ldg(reinterpret_cast<const float2*>(a.co + off + o * plane));
    }
    if (h0) {
      r.hgy[o].x = __ldg(a.gy + off + o * plane + hoff);
      if (HAS_GCO) r.hgco[o].x = __ldg(a.gco + off + o * plane + hoff);
      r.hy[o].x = __ldg(a.y + off + o * plane + hoff);
      r.hco[o].x = __ldg(a.co + off + o * plane + hoff);
    }
    if (h1) {
      r.hgy[o].y = __ldg(a.gy + off + o * plane + hoff + 1);
      if (HAS_GCO) r.hgco[o].y = __ldg(a.gco + off + o * plane + hoff + 1);
      r.hy[o].y = __ldg(a.y + off + o * plane + hoff + 1);
      r.hco[o].y = __ldg(a.co + off + o * plane + hoff + 1);
    }
  }
}

template <int CI, bool HAS_GCO>
__global__ void __launch_bounds__(BS_WARPS * 32, 1) nconv_bwd_stream5(const BsArgs a) {
  constexpr int CO = B5_CO, KK = B5_K * B5_K, NW = CO * CI * KK, NPU = CO * KK + 2 * CO;
  __shared__ float s_w[NW], s_is[CO], s_b[CO];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int i = tid; i < NW; i += BS_WARPS * 32) s_w[i] = __ldg(a.w + i);
  __syncthreads();
  if (tid < CO) {
    float s = 0.f;
    for (int j = 0; j < CI * KK; ++j) s += s_w[tid * CI * KK + j];
    s_is[tid] = 1.f / s;
    s_b[tid] = __ldg(a.bias + tid);
  }
  __syncthreads();

  const long long unit = (long long)blockIdx.x * BS_WARPS + warp;
  if (unit >= a.units) return;
  const int ic = (int)(unit % CI);                           // input channel of this unit (fastest: shares g loads with its twin)
  const long long u2 = unit / CI;
  const int strip = (int)(u2 % a.nstrip);
  const int seg = (int)((u2 / a.nstrip) % a.nseg);
  const int b = (int)(u2 / ((long long)a.nseg * a.nstrip));
  const int r0 = seg * a.RS, x0 = strip * B5_COLS, col = x0 + B5_PX * lane;
  const int rows = min(a.RS, a.H - r0);
  const int W = a.W, H = a.H;
  const size_t plane = (size_t)H * W;
  const bool inb = col < W;
  const int hcol = (lane == 0) ? x0 - 2 : x0 + B5_COLS;      // first of the two halo columns
  const bool edge = (lane == 0 || lane == 31);
  const bool h0 = edge && hcol >= 0 && hcol < W, h1 = edge && hcol + 1 >= 0 && hcol + 1 < W;
  const int hoff = (lane == 0) ? -2 : B5_PX;
  float sden[CO], isum[CO], bo[CO];
#pragma unroll
  for (int o = 0; o < CO; ++o) { isum[o] = s_is[o]; sden[o] = 1.f / isum[o]; bo[o] = s_b[o]; }
  const float* wi = s_w + ic * KK;                            // W[o][ic][u][v] = wi[o * CI * KK + u * 5 + v]

  long long goff = ((long long)b * CO * H + (r0 - 2)) * W + col;       // next g row to load
  long long xoff = (((long long)b * CI + ic) * H + r0) * W + col;      // next x row to load
  long long ooff = xoff;                                               // next gx/gc row to store

  u64 win[B5_K][CO][B5_WIN];            // (gN, gD) windows of the last five output-space rows
  float accW[CO][KK], accB[CO], accT[CO];
#pragma unroll
  for (int o = 0; o < CO; ++o) {
    accB[o] = 0.f; accT[o] = 0.f;
#pragma unroll
    for (int t = 0; t < KK; ++t) accW[o][t] = 0.f;
  }

  B5RawG rg;
  float2 rxx = make_float2(0.f, 0.f), rxc = make_float2(0.f, 0.f);
  {
    const bool ok = r0 - 2 >= 0;
    b5_load_g<HAS_GCO>(rg, a, goff, plane, hoff, ok && inb, ok && h0, ok && h1);
    goff += W;
    if (inb) {                                               // row r0: consumed at step 4
      rxx = __ldg(reinterpret_cast<const float2*>(a.x + xoff));
      rxc = __ldg(reinterpret_cast<const float2*>(a.c + xoff));
    }
    xoff += W;
  }

  const int nsteps = rows + 4;
#pragma unroll 1
  for (int sb = 0; sb < nsteps; sb += B5_K) {
#pragma unroll
    for (int sl = 0; sl < B5_K; ++sl) {
      const int s = sb + sl;
      if (s < nsteps) {
        // ---- output-space row gr = r0-2+s -> packed (gN, gD) window with its two-column halos
        {
          const int gr = r0 - 2 + s;
          const bool count = (ic == 0) && (s >= 2) && (s < rows + 2);
#pragma unroll
          for (int o = 0; o < CO; ++o) {
            const float gy2[2] = {rg.gy[o].x, rg.gy[o].y}, gc2[2] = {rg.gco[o].x, rg.gco[o].y};
            const float y2[2] = {rg.y[o].x, rg.y[o].y}, c2[2] = {rg.co[o].x, rg.co[o].y};
#pragma unroll
            for (int p = 0; p < 2; ++p) {
              const float gn = gy2[p] * rcp_fast(fmaf(c2[p], sden[o], a.eps));
              const float gd = fmaf(gc2[p], isum[o], -gn * (y2[p] - bo[o]));
              win[sl][o][p + 2] = pack2f(gn, gd);
              if (count) { accB[o] += gy2[p]; accT[o] = fmaf(gc2[p], c2[p], accT[o]); }
            }
            const float hg2[2] = {rg.hgy[o].x, rg.hgy[o].y}, hc2[2] = {rg.hgco[o].x, rg.hgco[o].y};
            const float hy2[2] = {rg.hy[o].x, rg.hy[o].y}, hco2[2] = {rg.hco[o].x, rg.hco[o].y};
            u64 hp[2];
#pragma unroll
            for (int p = 0; p < 2; ++p) {
              const float hgn = hg2[p] * rcp_fast(fmaf(hco2[p], sden[o], a.eps));
              hp[p] = pack2f(hgn, fmaf(hc2[p], isum[o], -hgn * (hy2[p] - bo[o])));
            }
            const u64 l0 = shfl_up64(win[sl][o][2]), l1 = shfl_up64(win[sl][o][3]);
            const u64 q0 = shfl_down64(win[sl][o][2]), q1 = shfl_down64(win[sl][o][3]);
            win[sl][o][0] = (lane == 0) ? hp[0] : l0;
            win[sl][o][1] = (lane == 0) ? hp[1] : l1;
            win[sl][o][4] = (lane == 31) ? hp[0] : q0;
            win[sl][o][5] = (lane == 31) ? hp[1] : q1;
          }
          const int gn_row = gr + 1;
          const bool ok = (s + 1 < nsteps) && gn_row >= 0 && gn_row < H;
          b5_load_g<HAS_GCO>(rg, a, goff, plane, hoff, ok && inb, ok && h0, ok && h1);
          goff += W;
        }
        // ---- input row q = r0+s-4: data gradient and weight-gradient accumulation for channel ic
        if (s >= 4) {
          const float xo[2] = {rxx.x, rxx.y}, cv[2] = {rxc.x, rxc.y};
          rxx = make_float2(0.f, 0.f); rxc = make_float2(0.f, 0.f);
          if (inb && (s + 1 < nsteps)) {
            rxx = __ldg(reinterpret_cast<const float2*>(a.x + xoff));
            rxc = __ldg(reinterpret_cast<const float2*>(a.c + xoff));
          }
          xoff += W;
          u64 A[2] = {0ull, 0ull};
          const float xc[2] = {xo[0] * cv[0], xo[1] * cv[1]};
#pragma unroll
          for (int o = 0; o < CO; ++o)
#pragma unroll
            for (int u = 0; u < B5_K; ++u) {
              const int slot = (sl - u + B5_K) % B5_K;               // output-space row q+2-u
#pragma unroll
              for (int v = 0; v < B5_K; ++v) {
                const float wv = wi[o * CI * KK + u * B5_K + v];
                float acc = accW[o][u * B5_K + v];
#pragma unroll
                for (int p = 0; p < 2; ++p) {
                  const u64 g2 = win[slot][o][p + 4 - v];
                  ffma2(A[p], g2, wv);
                  acc = fmaf(lo32f(g2), xc[p], acc);
                  acc = fmaf(hi32f(g2), cv[p], acc);
                }
                accW[o][u * B5_K + v] = acc;
              }
            }
          if (inb) {
            __stcs(reinterpret_cast<float2*>(a.gx + ooff), make_float2(lo32f(A[0]) * cv[0], lo32f(A[1]) * cv[1]));
            __stcs(reinterpret_cast<float2*>(a.gc + ooff),
                   make_float2(fmaf(lo32f(A[0]), xo[0], hi32f(A[0])), fmaf(lo32f(A[1]), xo[1], hi32f(A[1]))));
          }
          ooff += W;
        }
      }
    }
  }

  float* dst = a.partial + (size_t)unit * NPU;              // [o][tap] of channel ic, then gb[o], T[o] (channel-0 units)
#pragma unroll
  for (int o = 0; o < CO; ++o) {
#pragma unroll
    for (int t = 0; t < KK; ++t) {
      const float v = warp_sum(accW[o][t]);
      if (lane == 0) dst[o * KK + t] = v;
    }
    const float vb = warp_sum(accB[o]), vt = warp_sum(accT[o]);
    if (lane == 0) { dst[CO * KK + o] = vb; dst[CO * KK + CO + o] = vt; }
  }
}

// gW[o][i][t] = sum over the units of channel i - T[o] / sum(W[o]);  gb[o], T[o] from the channel-0 units.
template <int CI>
__global__ void nconv_bwd_stream5_final(const float* __restrict__ partial, const float* __restrict__ w, float* __restrict__ gw,
                                        float* __restrict__ gb, int units) {
  constexpr int CO = B5_CO, KK = B5_K * B5_K, NW = CO * CI * KK, NPU = CO * KK + 2 * CO;
  const int lane = threadIdx.x & 31;
  const int idx = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (idx >= NW + CO) return;
  if (idx < NW) {
    const int o = idx / (CI * KK), i = (idx / KK) % CI, t = idx % KK;
    float s = 0.f, T = 0.f, ws = 0.f;
    for (int u = i + CI * lane; u < units; u += CI * 32) s += __ldg(partial + (size_t)u * NPU + o * KK + t);
    for (int u = CI * lane; u < units; u += CI * 32) T += __ldg(partial + (size_t)u * NPU + CO * KK + CO + o);
    for (int j = lane; j < CI * KK; j += 32) ws += __ldg(w + o * CI * KK + j);
    s = warp_sum(s); T = warp_sum(T); ws = warp_sum(ws);
    if (lane == 0) gw[idx] = s - T / ws;
  } else if (gb) {
    const int o = idx - NW;
    float s = 0.f;
    for (int u = CI * lane; u < units; u += CI * 32) s += __ldg(partial + (size_t)u * NPU + CO * KK + o);
    s = warp_sum(s);
    if (lane == 0) gb[o] = s;
  }
}

template <int CI>
static int launch_bwd_stream5(BsArgs a, float* gw, float* gb, cudaStream_t st) {
  constexpr int CO = B5_CO, KK = B5_K * B5_K, NW = CO * CI * KK, NPU = CO * KK + 2 * CO;
  int dev = 0, sms = 0;
  NCONV_RETURN_IF_ERR(cudaGetDevice(&dev));
  static int sm_cache[NCONV_MAX_DEVICES] = {};
  if (dev < NCONV_MAX_DEVICES) sms = sm_cache[dev];
  if (sms == 0) {
    NCONV_RETURN_IF_ERR(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    if (dev < NCONV_MAX_DEVICES) sm_cache[dev] = sms;
  }
  const int nstrip = ceil_div(a.W, B5_COLS);
  const double cap = (double)sms * BS_WARPS;
  int best_rs = 0; double best = 1e300;
  for (int rs = 4; rs <= 160; ++rs) {
    const double units = (double)a.B * nstrip * ceil_div(a.H, rs) * CI;
    if (units > BS_MAX_UNITS) continue;
    const double cost = ceil(units / cap) * (rs + 4);
    if (cost < best - 1e-9) { best = cost; best_rs = rs; }
  }
  if (best_rs == 0) return NCONV_ERR_UNSUPPORTED;
  if (const char* e = getenv("NCONV_B200_BWD_STREAM_RS")) { int v = atoi(e); if (v > 0) best_rs = v; }
  a.RS = best_rs; a.nstrip = nstrip; a.nseg = ceil_div(a.H, best_rs);
  a.units = (long long)a.B * a.nstrip * a.nseg * CI;
  if (a.units > BS_MAX_UNITS) return NCONV_ERR_UNSUPPORTED;
  const unsigned blocks = (unsigned)((a.units + BS_WARPS - 1) / BS_WARPS);
  const double px = (double)a.B * a.H * a.W;
  char label[64];
  snprintf(label, sizeof(label), "nconv_bwd_stream %d->%d k5", CI, CO);
  {
    LaunchScope ls(st, label, 8.0 * CI * CO * KK * px, 4.0 * (4.0 * CO + 4.0 * CI) * px);
    if (a.gco) nconv_bwd_stream5<CI, true><<<blocks, BS_WARPS * 32, 0, st>>>(a);
    else nconv_bwd_stream5<CI, false><<<blocks, BS_WARPS * 32, 0, st>>>(a);
  }
  NCONV_RETURN_IF_ERR(cudaGetLastError());
  {
    LaunchScope ls(st, "nconv_bwd_stream_final", 0.0, 4.0 * (double)a.units * NPU);
    nconv_bwd_stream5_final<CI><<<ceil_div((NW + CO) * 32, 128), 128, 0, st>>>(a.partial, a.w, gw, gb, (int)a.units);
  }
  return (int)cudaGetLastError();
}

}  // namespace nconv
