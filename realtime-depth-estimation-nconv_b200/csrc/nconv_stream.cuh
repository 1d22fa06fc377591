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

// Per-lane addressing state: everything that does not change along the walk is computed once.
struct StLane {
  const float* px; const float* pc;   // own 4 px of (channel 0, first input row); advanced by W per row
  int hoff;                           // halo offset in floats relative to px (edge lanes)
  bool inb, hload;                    // own px inside the image / this lane fetches a halo
};

template <int PAD, bool THRESH>
__device__ __forceinline__ void st_load(ChanRow<PAD>& r, const StLane& L, size_t off, bool rowok) {
  const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
  r.x = z; r.c = z;
#pragma unroll
  for (int j = 0; j < PAD; ++j) { r.hx[j] = 0.f; r.hc[j] = 0.f; }
  if (rowok && L.inb) {
    r.x = __ldg(reinterpret_cast<const float4*>(L.px + off));
    if (!THRESH) r.c = __ldg(reinterpret_cast<const float4*>(L.pc + off));
  }
  if (rowok && L.hload) {
    if (PAD == 2) {
      const float2 hx = __ldg(reinterpret_cast<const float2*>(L.px + off + L.hoff));
      r.hx[0] = hx.x; r.hx[PAD - 1] = hx.y;
      if (!THRESH) { const float2 hc = __ldg(reinterpret_cast<const float2*>(L.pc + off + L.hoff)); r.hc[0] = hc.x; r.hc[PAD - 1] = hc.y; }
    } else {
      r.hx[0] = __ldg(L.px + off + L.hoff);
      if (!THRESH) r.hc[0] = __ldg(L.pc + off + L.hoff);
    }
  }
}

template <int CI, int CO, int K, int PF, bool THRESH>
__global__ void __launch_bounds__(ST_WARPS * 32) nconv_stream(const StreamArgs a) {
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
  const int seg = (int)(unit % a.nseg);
  const int strip = (int)((unit / a.nseg) % a.nstrip);
  const int b = (int)(unit / ((long long)a.nseg * a.nstrip));
  const int oy0 = seg * a.RS, x0 = strip * ST_COLS, col = x0 + ST_PX * lane;
  const int rows = min(a.RS, a.H - oy0);
  const int T = rows + 2 * PAD;                   // input rows this unit walks (some may lie outside the image)
  const int W = a.W;
  const size_t plane = (size_t)a.H * W;

  StLane L;
  {
    // pointer of (channel 0, input row oy0-PAD); rows outside the image are never dereferenced
    const long long first = ((long long)b * CI * a.H + (oy0 - PAD)) * W + col;
    L.px = a.x + first; L.pc = THRESH ? a.x + first : a.c + first;
    L.inb = col < W;
    const int hcol = (lane == 0) ? x0 - PAD : x0 + ST_COLS;
    L.hload = (lane == 0 || lane == 31) && hcol >= 0 && hcol < W;
    L.hoff = (lane == 0) ? -PAD : ST_PX;
  }
  const int t_lo = max(0, PAD - oy0), t_hi = min(T, a.H + PAD - oy0);   // input rows inside the image

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
    st_load<PAD, THRESH>(ring[i], L, (size_t)t * W + ci * plane, t >= t_lo && t < t_hi);
  }

  float* ypt = a.y + ((size_t)b * CO * a.H + oy0) * W + col;      // output row 0 of channel 0 (own px)
  float* cpt = a.co + ((size_t)b * CO * a.H + oy0) * W + col;

#pragma unroll 1
  for (int tb = 0; tb < T; tb += U / CI) {
#pragma unroll
    for (int sl = 0; sl < U; ++sl) {
      const int tl = sl / CI, ci = sl % CI, ph = tl % K;     // static after unrolling (U % (K*CI) == 0)
      const int t = tb + tl;
      if (t < T) {
        const bool rowok = t >= t_lo && t < t_hi;
        if (rowok) {
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
            const float lx = __shfl_up_sync(0xffffffffu, xw[ST_PX + j], 1);
            const float lc = __shfl_up_sync(0xffffffffu, cw[ST_PX + j], 1);
            const float rx = __shfl_down_sync(0xffffffffu, xw[PAD + j], 1);
            const float rc = __shfl_down_sync(0xffffffffu, cw[PAD + j], 1);
            xw[j] = (lane == 0) ? hxc : lx;               cw[j] = (lane == 0) ? hc : lc;
            xw[PAD + ST_PX + j] = (lane == 31) ? hxc : rx; cw[PAD + ST_PX + j] = (lane == 31) ? hc : rc;
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
        } else if (ci == 0) {
          // input row outside the image: the output row that would start here starts from zero
#pragma unroll
          for (int o = 0; o < CO_P; ++o)
#pragma unroll
            for (int p = 0; p < ST_PX; ++p) { accN[ph][o][p] = 0ull; accD[ph][o][p] = 0ull; }
        }
        // ---- refill the ring slot PF steps ahead
        {
          const int tn = tb + (sl + PF) / CI, cn = (sl + PF) % CI;
          st_load<PAD, THRESH>(ring[sl % PF], L, (size_t)tn * W + cn * plane, tn >= t_lo && tn < t_hi);
        }
        if (ci == CI - 1) {
          // ---- output row t-(K-1) has now seen all of its K input rows
          const int slot = (ph + 1) % K;
          const int ro = t - (K - 1);
          if (ro >= 0 && L.inb) {
#pragma unroll
            for (int o = 0; o < CO; ++o) {
              float yv[ST_PX], cvv[ST_PX];
              const float bo = s_bias[o], rs = s_rs[o];
#pragma unroll
              for (int p = 0; p < ST_PX; ++p) {
                const float N = (o & 1) ? hi32f(accN[slot][o >> 1][p]) : lo32f(accN[slot][o >> 1][p]);
                const float D = (o & 1) ? hi32f(accD[slot][o >> 1][p]) : lo32f(accD[slot][o >> 1][p]);
                yv[p] = __fdividef(N, D + a.eps) + bo;
                cvv[p] = D * rs;
              }
              const size_t off = o * plane + (size_t)ro * W;
              __stcs(reinterpret_cast<float4*>(ypt + off), make_float4(yv[0], yv[1], yv[2], yv[3]));
              __stcs(reinterpret_cast<float4*>(cpt + off), make_float4(cvv[0], cvv[1], cvv[2], cvv[3]));
            }
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
