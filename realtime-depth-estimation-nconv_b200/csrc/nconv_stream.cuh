// Streaming NConv2d forward for the HBM-bound small-channel shapes (Ci*Co <= 16, K in {3,5}) —
// /root/reference/models/step1.py:116-149 (same arithmetic as nconv_tiled.cuh, other schedule).
//
// These shapes move 24..64 B per pixel for 18..288 FMA pairs: the roofline is HBM, so the kernel is
// built around keeping loads in flight, not around operand reuse through shared memory:
//   * one WARP owns a strip of 128 columns (4 px per lane) x RS output rows of one image and walks
//     it top to bottom; there is no shared-memory staging of activations and no block barrier in
//     the loop (weights sit in shared memory, read as broadcast LDS);
//   * an input row is loaded ONCE per strip as one 128-bit coalesced load per lane per plane, PF
//     rows ahead of its use (register prefetch ring), the column halo comes from the neighbouring
//     lanes by shuffle (the two edge lanes fetch theirs from global: same lines, L1/L2 hits);
//   * the K output rows an input row contributes to are K rolling accumulator sets in registers
//     (input-stationary): row t updates sets (t-u) mod K for tap rows u = 0..K-1, after which set
//     (t-K+1) mod K is complete, goes through the epilogue (y = N/(D+eps)+b, cout = D/sum(W)),
//     is stored with 128-bit stores and restarts at zero.  The loop is unrolled by K so every
//     set index is a compile-time constant;
//   * FMAs are packed (FFMA2: two output channels per instruction, common.cuh).
// Vertical halo: a strip segment re-reads K-1 rows of its neighbour ((K-1)/RS extra, L2 hits).
#pragma once
#include "common.cuh"
#include "nconv_b200.h"
#include "launch_count.h"
#include <cstdio>
#include <cstdlib>
#include <cmath>

namespace nconv {

constexpr int ST_WARPS = 8;          // warps per CTA (independent units)
constexpr int ST_PX = 4;             // pixels per lane
constexpr int ST_COLS = 32 * ST_PX;  // strip width

struct StreamArgs {
  const float* x; const float* c; const float* w; const float* bias;
  float* y; float* co;
  int B, H, W;            // input = output size (pad = K/2)
  int RS;                 // output rows per unit
  int nstrip, nseg;       // strips per row, row segments per image
  long long units;        // B * nstrip * nseg
  float eps, thr;
};

template <int CI, int CO, int K>
struct StreamSmem {
  static constexpr int W_FLOATS = CI * K * K * CO;   // [ci][u][v][o]
  static constexpr size_t BYTES = sizeof(float) * (W_FLOATS + 2 * CO);
};

// One prefetched (row, channel) of a lane: 4 own px of both planes + the edge-lane halo.
template <int PAD>
struct ChanRow {
  float4 x, c;
  float hx[PAD], hc[PAD];
};

__host__ __device__ constexpr int st_gcd(int a, int b) { return b == 0 ? a : st_gcd(b, a % b); }
__host__ __device__ constexpr int st_lcm(int a, int b) { return a / st_gcd(a, b) * b; }

// acc = w * s (first tap row of an output row: no separate zeroing of the accumulator set)
__device__ __forceinline__ void fmul2(u64& acc, u64 w, float s) {
  asm("{\n\t.reg .b64 t;\n\tmov.b64 t, {%2, %2};\n\tmul.rn.f32x2 %0, %1, t;\n\t}" : "=l"(acc) : "l"(w), "f"(s));
}

// 1/v for v >= eps > 0 (no zero / denormal / infinity handling needed): one MUFU.RCP, 1 ulp.
__device__ __forceinline__ float rcp_fast(float v) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(v));
  return r;
}

// One prefetch: rows outside the image and columns outside the row stay zero (zero padding), so the
// compute step needs no branch — a padded row simply adds zeros.  Pointers may point outside the
// tensor when the predicate is false; they are never dereferenced then.
template <int PAD, bool THRESH>
__device__ __forceinline__ void st_load(ChanRow<PAD>& r, const float* px, const float* pc, int hoff,
                                        bool own, bool halo) {
  const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
  r.x = z; r.c = z;
#pragma unroll
  for (int j = 0; j < PAD; ++j) { r.hx[j] = 0.f; r.hc[j] = 0.f; }
  if (own) {
    r.x = __ldg(reinterpret_cast<const float4*>(px));
    if (!THRESH) r.c = __ldg(reinterpret_cast<const float4*>(pc));
  }
  if (halo) {
    if (PAD == 2) {
      const float2 hx = __ldg(reinterpret_cast<const float2*>(px + hoff));
      r.hx[0] = hx.x; r.hx[PAD - 1] = hx.y;
      if (!THRESH) { const float2 hc = __ldg(reinterpret_cast<const float2*>(pc + hoff)); r.hc[0] = hc.x; r.hc[PAD - 1] = hc.y; }
    } else {
      r.hx[0] = __ldg(px + hoff);
      if (!THRESH) r.hc[0] = __ldg(pc + hoff);
    }
  }
}

template <int CI, int CO, int K, int PF, bool THRESH>
__global__ void __launch_bounds__(ST_WARPS * 32, (K == 3 && CI * CO <= 4 && PF <= 3) ? 2 : 1) nconv_stream(const StreamArgs a) {
  constexpr int PAD = K / 2, CO_P = CO / 2, WIN = ST_PX + 2 * PAD;
  constexpr int U = st_lcm(K * CI, PF);          // unroll: accumulator-set, channel and ring indices all static
  static_assert(CO % 2 == 0 && (K == 3 || K == 5), "stream kernel: even CO, K in {3,5}");
  using SM = StreamSmem<CI, CO, K>;
  extern __shared__ __align__(16) float st_smem[];
  float* s_w = st_smem;                 // [ci][u][v][o]
  float* s_bias = s_w + SM::W_FLOATS;   // [o]
  float* s_rs = s_bias + CO;            // [o] 1 / sum(W[o])

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int idx = tid; idx < SM::W_FLOATS; idx += ST_WARPS * 32) {
    int o = idx % CO, r = idx / CO;
    int v = r % K; r /= K;
    int u = r % K; r /= K;
    s_w[idx] = __ldg(&a.w[(((size_t)o * CI + r) * K + u) * K + v]);
  }
  __syncthreads();
  for (int o = warp; o < CO; o += ST_WARPS) {
    float s = 0.f;
    for (int j = lane; j < CI * K * K; j += 32) s += s_w[j * CO + o];
    s = warp_sum(s);
    if (lane == 0) { s_rs[o] = 1.f / s; s_bias[o] = __ldg(&a.bias[o]); }
  }
  __syncthreads();

  const long long unit = (long long)blockIdx.x * ST_WARPS + warp;
  if (unit >= a.units) return;
  // strip index fastest: the warps of a CTA (and neighbouring CTAs) walk adjacent strips of the SAME rows, so
  // what is in flight at any moment is whole contiguous image rows (DRAM page locality), not 512 B pieces
  const int strip = (int)(unit % a.nstrip);
  const int seg = (int)((unit / a.nstrip) % a.nseg);
  const int b = (int)(unit / ((long long)a.nseg * a.nstrip));
  const int oy0 = seg * a.RS, x0 = strip * ST_COLS, col = x0 + ST_PX * lane;
  const int rows = min(a.RS, a.H - oy0);
  const int T = rows + 2 * PAD;                   // input rows this unit walks (some may lie outside the image)
  const int W = a.W;
  const size_t plane = (size_t)a.H * W;

  // ---- per-lane state that does not change along the walk
  const bool inb = col < W;
  const int hcol = (lane == 0) ? x0 - PAD : x0 + ST_COLS;
  const bool hlane = (lane == 0 || lane == 31) && hcol >= 0 && hcol < W;
  const int hoff = (lane == 0) ? -PAD : ST_PX;
  // input rows inside the image are t in [t_lo, t_hi): one unsigned compare per prefetch
  const int t_lo = max(0, PAD - oy0);
  const unsigned t_n = (unsigned)(min(T, a.H + PAD - oy0) - t_lo);
  // load cursor: (row of the next prefetch, channel 0), own 4 px
  const long long first = ((long long)b * CI * a.H + (oy0 - PAD)) * W + col;
  const float* lx = a.x + first;
  const float* lc = THRESH ? lx : a.c + first;
  // store cursor: output row 0 of channel 0 (own px); advanced by W per finished row
  float* yp = a.y + ((size_t)b * CO * a.H + oy0) * W + col;
  float* cp = a.co + ((size_t)b * CO * a.H + oy0) * W + col;

  u64 accN[K][CO_P][ST_PX], accD[K][CO_P][ST_PX];
#pragma unroll
  for (int k = 0; k < K; ++k)
#pragma unroll
    for (int o = 0; o < CO_P; ++o)
#pragma unroll
      for (int p = 0; p < ST_PX; ++p) { accN[k][o][p] = 0ull; accD[k][o][p] = 0ull; }

  ChanRow<PAD> ring[PF];
#pragma unroll
  for (int i = 0; i < PF; ++i) {
    const int t = i / CI, ci = i % CI;
    const bool ok = (unsigned)(t - t_lo) < t_n;
    st_load<PAD, THRESH>(ring[i], lx + ci * plane, lc + ci * plane, hoff, ok && inb, ok && hlane);
    if (ci == CI - 1) { lx += W; lc += W; }
  }
  int tl_next = PF / CI;      // row index of the load cursor

#pragma unroll 1
  for (int tb = 0; tb < T; tb += U / CI) {
#pragma unroll
    for (int sl = 0; sl < U; ++sl) {
      const int tl = sl / CI, ci = sl % CI, ph = tl % K;     // static after unrolling (U % (K*CI) == 0)
      const int t = tb + tl;
      if (t < T) {
        {
          // ---- window of this (row, channel): own px, neighbours' px by shuffle, strip-edge halo from the ring
          const ChanRow<PAD>& r = ring[sl % PF];
          float xw[WIN], cw[WIN];
          const float xo[4] = {r.x.x, r.x.y, r.x.z, r.x.w};
          float cv[4] = {r.c.x, r.c.y, r.c.z, r.c.w};
#pragma unroll
          for (int p = 0; p < ST_PX; ++p) {
            if (THRESH) cv[p] = (xo[p] > a.thr) ? 1.f : 0.f;
            cw[PAD + p] = cv[p];
            xw[PAD + p] = xo[p] * cv[p];
          }
#pragma unroll
          for (int j = 0; j < PAD; ++j) {
            float hc = r.hc[j];
            if (THRESH) hc = (r.hx[j] > a.thr) ? 1.f : 0.f;
            const float hxc = r.hx[j] * hc;
            const float lxv = __shfl_up_sync(0xffffffffu, xw[ST_PX + j], 1);
            const float lcv = __shfl_up_sync(0xffffffffu, cw[ST_PX + j], 1);
            const float rxv = __shfl_down_sync(0xffffffffu, xw[PAD + j], 1);
            const float rcv = __shfl_down_sync(0xffffffffu, cw[PAD + j], 1);
            xw[j] = (lane == 0) ? hxc : lxv;               cw[j] = (lane == 0) ? hc : lcv;
            xw[PAD + ST_PX + j] = (lane == 31) ? hxc : rxv; cw[PAD + ST_PX + j] = (lane == 31) ? hc : rcv;
          }
          const float* wci = s_w + ci * K * K * CO;
#pragma unroll
          for (int u = 0; u < K; ++u) {
            const int slot = (ph - u + K) % K;
#pragma unroll
            for (int v = 0; v < K; ++v) {
              u64 w2[CO_P];
              if constexpr (CO % 4 == 0) {
#pragma unroll
                for (int o = 0; o < CO_P; o += 2) {
                  const ulonglong2 tt = *reinterpret_cast<const ulonglong2*>(wci + (u * K + v) * CO + 2 * o);
                  w2[o] = tt.x; w2[o + 1] = tt.y;
                }
              } else {
#pragma unroll
                for (int o = 0; o < CO_P; ++o) w2[o] = *reinterpret_cast<const u64*>(wci + (u * K + v) * CO + 2 * o);
              }
#pragma unroll
              for (int o = 0; o < CO_P; ++o)
#pragma unroll
                for (int p = 0; p < ST_PX; ++p) {
                  if (u == 0 && v == 0 && ci == 0) {          // first contribution to a fresh output row
                    fmul2(accN[slot][o][p], w2[o], xw[p + v]);
                    fmul2(accD[slot][o][p], w2[o], cw[p + v]);
                  } else {
                    ffma2(accN[slot][o][p], w2[o], xw[p + v]);
                    ffma2(accD[slot][o][p], w2[o], cw[p + v]);
                  }
                }
            }
          }
        }
        // ---- refill the ring slot PF steps ahead (cursor arithmetic is static apart from the row test)
        {
          const int cn = (sl + PF) % CI;
          const bool ok = (unsigned)(tl_next - t_lo) < t_n;
          st_load<PAD, THRESH>(ring[sl % PF], lx + cn * plane, lc + cn * plane, hoff, ok && inb, ok && hlane);
          if (cn == CI - 1) { lx += W; lc += W; ++tl_next; }
        }
        if (ci == CI - 1) {
          // ---- output row t-(K-1) has now seen all of its K input rows
          const int slot = (ph + 1) % K;
          if (t >= K - 1) {
            if (inb) {
#pragma unroll
              for (int o = 0; o < CO; ++o) {
                float yv[ST_PX], cvv[ST_PX];
                const float bo = s_bias[o], rs = s_rs[o];
#pragma unroll
                for (int p = 0; p < ST_PX; ++p) {
                  const float N = (o & 1) ? hi32f(accN[slot][o >> 1][p]) : lo32f(accN[slot][o >> 1][p]);
                  const float D = (o & 1) ? hi32f(accD[slot][o >> 1][p]) : lo32f(accD[slot][o >> 1][p]);
                  yv[p] = fmaf(N, rcp_fast(D + a.eps), bo);
                  cvv[p] = D * rs;
                }
                __stcs(reinterpret_cast<float4*>(yp + o * plane), make_float4(yv[0], yv[1], yv[2], yv[3]));
                __stcs(reinterpret_cast<float4*>(cp + o * plane), make_float4(cvv[0], cvv[1], cvv[2], cvv[3]));
              }
            }
            yp += W; cp += W;
          }
        }
      }
    }
  }
}

template <int CI, int CO, int K, int PF, bool THRESH>
static int launch_stream(const float* x, const float* c, const float* w, const float* bias, float* y, float* co,
                         int B, int H, int W, float eps, float thr, cudaStream_t st) {
  using SM = StreamSmem<CI, CO, K>;
  auto kern = nconv_stream<CI, CO, K, PF, THRESH>;
  static int occ_cache[NCONV_MAX_DEVICES] = {};
  static int sm_cache[NCONV_MAX_DEVICES] = {};
  int dev = 0;
  NCONV_RETURN_IF_ERR(cudaGetDevice(&dev));
  int occ = dev < NCONV_MAX_DEVICES ? occ_cache[dev] : 0, sms = dev < NCONV_MAX_DEVICES ? sm_cache[dev] : 0;
  if (occ == 0) {
    NCONV_RETURN_IF_ERR(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, ST_WARPS * 32, SM::BYTES));
    NCONV_RETURN_IF_ERR(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    if (occ < 1) occ = 1;
    if (dev < NCONV_MAX_DEVICES) { occ_cache[dev] = occ; sm_cache[dev] = sms; }
  }
  // Rows per unit: the grid should fill whole waves of resident warps; a unit costs RS + K - 1 rows.
  const int nstrip = ceil_div(W, ST_COLS);
  const double cap = (double)sms * occ * ST_WARPS;
  int best_rs = 16; double best = 1e300;
  for (int rs = 6; rs <= 96; ++rs) {
    const double units = (double)B * nstrip * ceil_div(H, rs);
    const double waves = ceil(units / cap);
    const double cost = waves * (rs + K - 1) * (1.0 + 0.01 * (K - 1) * 16.0 / rs);  // mild preference for less halo re-read
    if (cost < best - 1e-9) { best = cost; best_rs = rs; }
  }
  if (const char* e = getenv("NCONV_B200_STREAM_RS")) { int v = atoi(e); if (v > 0) best_rs = v; }
  StreamArgs a = {};
  a.x = x; a.c = c; a.w = w; a.bias = bias; a.y = y; a.co = co;
  a.B = B; a.H = H; a.W = W; a.RS = best_rs; a.nstrip = nstrip; a.nseg = ceil_div(H, best_rs);
  a.units = (long long)B * a.nstrip * a.nseg; a.eps = eps; a.thr = thr;
  const long long blocks = (a.units + ST_WARPS - 1) / ST_WARPS;
  const double px = (double)B * H * W;
  char label[64];
  snprintf(label, sizeof(label), "nconv_stream %d->%d k%d f%d", CI, CO, K, THRESH ? 1 : 0);
  {
    LaunchScope ls(st, label, 4.0 * CI * CO * K * K * px, 4.0 * ((THRESH ? 1.0 : 2.0) * CI * px + 2.0 * CO * px));
    kern<<<(unsigned)blocks, ST_WARPS * 32, SM::BYTES, st>>>(a);
  }
  return (int)cudaGetLastError();
}

}  // namespace nconv
