// Persistent, software-pipelined variant of the tiled NConv2d forward (nconv_tiled.cuh) for the FMA-bound DNET
// layers (/root/reference/models/step1.py:38-49: 8->8 k5, [8|8]->8 k3, 1->8 k5 and the nconv6+nconv7 tail).
//
// nconv_tiled spends a third of every CTA's life outside the FFMA2 loop (ncu: fma pipe 65 % active, stall samples
// 17 % long-scoreboard in the staging loop, weights prologue once per 16x64 tile).  Here
//   * a CTA is persistent: the weight transpose / sums / bias prologue runs once, then the CTA walks tiles
//     t = blockIdx.x, blockIdx.x + gridDim.x, ...;
//   * staging is asynchronous: the raw x and c quads of chunk q+1 are copied global -> shared with cp.async
//     (zero-filled outside the image through the src-size operand) into the second buffer while chunk q is being
//     accumulated, across tile boundaries too, so the epilogue's stores overlap the next tile's loads;
//   * when a chunk is due, each thread waits for ITS OWN copies, fixes its slots up in place (x -> x*c; threshold
//     -> c for nconv1; 2x nearest expansion for the low-resolution source) and one __syncthreads publishes it.
// Same arithmetic, tile shape, accumulation order and epilogues as nconv_tiled: results are bit-identical to it.
#pragma once
#include "nconv_tiled.cuh"

namespace nconv {

template <int CA, int CB, int CO, int K>
struct PipeSmem {
  static constexpr int CI = CA + CB;
  static constexpr int CCH = CI < 4 ? CI : 4;
  static constexpr int NCHUNK = (CI + CCH - 1) / CCH;
  static constexpr int RH = TH + K - 1;
  static constexpr int W_FLOATS = ((CI * K * K * CO) + 3) & ~3;
  static constexpr int TILE_FLOATS = CCH * RH * SW;
  static constexpr size_t BYTES = sizeof(float) * (W_FLOATS + 16 + 4 * TILE_FLOATS);
};

// One CTA produces CO (even, <= 8) of the layer's CO_ALL output channels: group blockIdx.y.  FLAGS: F_THRESH |
// F_POOL | F_TAIL as in nconv_tiled.
template <int CA, int CB, int CO_ALL, int CO, int K, int PAD, int FLAGS>
__global__ void __launch_bounds__(NTHREADS, 2) nconv_pipe(const TiledArgs a, int tiles_x, int tiles_y, int ntiles) {
  using SM = PipeSmem<CA, CB, CO, K>;
  constexpr int CI = SM::CI, CCH = SM::CCH, NCHUNK = SM::NCHUNK, RH = SM::RH;
  constexpr bool THRESH = (FLAGS & NCONV_B200_F_THRESH) != 0;
  constexpr bool POOL = (FLAGS & NCONV_B200_F_POOL) != 0;
  constexpr bool TAIL = (FLAGS & NCONV_B200_F_TAIL) != 0;
  constexpr int OFF = (4 - (PAD & 3)) & 3;
  constexpr int NV = (OFF + PX + K - 1 + 3) / 4;
  constexpr int CO_P = CO / 2;
  constexpr int SLOTS = CCH * RH * (SW / 4);
  constexpr int NIT = (SLOTS + NTHREADS - 1) / NTHREADS;
  static_assert(CO % 2 == 0 && CO <= 8 && CO_ALL % CO == 0, "nconv_pipe: CO must be even, <= 8 and divide CO_ALL");
  static_assert(!TAIL || CO_ALL == CO, "tail needs all channels in one CTA");
  const int cog = blockIdx.y;
  static_assert(CB == 0 || CA % CCH == 0, "nconv_pipe: a chunk must come from one source");
  static_assert(CI % CCH == 0, "nconv_pipe: whole chunks only");
  static_assert(4 * NV + 4 * (TW / PX - 1) <= SW, "staged row too narrow");

  extern __shared__ __align__(16) float smem[];
  float* s_w = smem;                         // [CI][K][K][CO]
  float* s_sum = s_w + SM::W_FLOATS;         // [CO] 1 / sum of weights
  float* s_bias = s_sum + 8;                 // [CO]
  float* s_buf = s_bias + 8;                 // [2 buffers][x | c][CCH][RH][SW]

  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;

  // ---- chunk issue: asynchronous copies of the raw quads of (tile t, chunk ch) into buffer `bufi`
  auto issue = [&](int t, int ch, int bufi) {
    const int bx = t % tiles_x, by = (t / tiles_x) % tiles_y, b = t / (tiles_x * tiles_y);
    const int gx0a = bx * TW - PAD - OFF, gy0 = by * TH - PAD;
    const int c0 = ch * CCH;
    float* dx = s_buf + (size_t)bufi * 2 * SM::TILE_FLOATS;
    float* dc = dx + SM::TILE_FLOATS;
    const bool from_a = (CB == 0) || (c0 < CA);
#pragma unroll
    for (int it = 0; it < NIT; ++it) {
      const int idx = tid + it * NTHREADS;
      if (idx < SLOTS) {
        const int j = idx % (SW / 4), r = (idx / (SW / 4)) % RH, cc = idx / ((SW / 4) * RH);
        const int cs = c0 + cc, gy = gy0 + r, gx = gx0a + 4 * j;
        const bool ok = gy >= 0 && gy < a.H && gx >= 0 && gx + 3 < a.W;
        const int so = 4 * idx;               // == (cc*RH + r)*SW + 4*j
        if (from_a) {
          const size_t base = ok ? ((size_t)(b * CA + cs) * a.H + gy) * a.W + gx : 0;
          cp_async16(dx + so, a.xa + base, ok);
          if (!THRESH) cp_async16(dc + so, a.ca + base, ok);
        } else {
          const size_t base = ok ? ((size_t)(b * CB + (cs - CA)) * a.HB + (gy >> 1)) * a.WB + (gx >> 1) : 0;
          cp_async8(dx + so, a.xb + base, ok);
          cp_async8(dc + so, a.cb + base, ok);
        }
      }
    }
  };
  // ---- in-place fix-up of this thread's own slots of buffer `bufi` holding chunk ch
  auto fixup = [&](int ch, int bufi) {
    float* dx = s_buf + (size_t)bufi * 2 * SM::TILE_FLOATS;
    float* dc = dx + SM::TILE_FLOATS;
    const bool from_a = (CB == 0) || (ch * CCH < CA);
#pragma unroll
    for (int it = 0; it < NIT; ++it) {
      const int idx = tid + it * NTHREADS;
      if (idx < SLOTS) {
        const int so = 4 * idx;               // slots are in memory order: (cc*RH + r)*SW + 4*j == 4*idx
        if (from_a) {
          float4 x4 = *reinterpret_cast<const float4*>(dx + so);
          float4 c4;
          if (THRESH) {
            // out-of-image quads were zero-filled: (0 > thr) is false for the thr >= 0 this path is used with
            c4 = make_float4(x4.x > a.thr ? 1.f : 0.f, x4.y > a.thr ? 1.f : 0.f, x4.z > a.thr ? 1.f : 0.f,
                             x4.w > a.thr ? 1.f : 0.f);
            *reinterpret_cast<float4*>(dc + so) = c4;
          } else {
            c4 = *reinterpret_cast<const float4*>(dc + so);
          }
          *reinterpret_cast<float4*>(dx + so) = make_float4(x4.x * c4.x, x4.y * c4.y, x4.z * c4.z, x4.w * c4.w);
        } else {
          const float2 x2 = *reinterpret_cast<const float2*>(dx + so);
          const float2 c2 = *reinterpret_cast<const float2*>(dc + so);
          const float p0 = x2.x * c2.x, p1 = x2.y * c2.y;
          *reinterpret_cast<float4*>(dx + so) = make_float4(p0, p0, p1, p1);
          *reinterpret_cast<float4*>(dc + so) = make_float4(c2.x, c2.x, c2.y, c2.y);
        }
      }
    }
  };

  int t = blockIdx.x;
  if (t < ntiles) issue(t, 0, 0);            // in flight during the weight prologue

  // ---- prologue (once per CTA): weights -> smem as [staged channel][u][v][o]; staged order is A then B
  for (int idx = tid; idx < CI * K * K * CO; idx += NTHREADS) {
    const int o = idx % CO;
    int r = idx / CO;
    const int v = r % K; r /= K;
    const int u = r % K; r /= K;
    const int cs = r;
    int ciw;
    if (CB == 0) ciw = cs;
    else if (a.a_first) ciw = cs;
    else ciw = (cs < CA) ? (CB + cs) : (cs - CA);
    s_w[idx] = __ldg(&a.w[(((size_t)(cog * CO + o) * CI + ciw) * K + u) * K + v]);
  }
  __syncthreads();
  {
    const int warp = tid >> 5, lane = tid & 31;
    for (int o = warp; o < CO; o += NTHREADS / 32) {
      float s = 0.f;
      for (int j = lane; j < CI * K * K; j += 32) s += s_w[j * CO + o];
      s = warp_sum(s);
      if (lane == 0) { s_sum[o] = 1.f / s; s_bias[o] = __ldg(&a.bias[cog * CO + o]); }
    }
  }
  // (s_sum / s_bias are first read in an epilogue, after at least one __syncthreads of the chunk loop)

  int q = 0;                                  // chunks consumed so far by this CTA
  for (; t < ntiles; t += gridDim.x) {
    const int bx = t % tiles_x, by = (t / tiles_x) % tiles_y, b = t / (tiles_x * tiles_y);
    const int ox0 = bx * TW, oy0 = by * TH;
    u64 accN2[CO_P][PX], accD2[CO_P][PX];
#pragma unroll
    for (int o = 0; o < CO_P; ++o)
#pragma unroll
      for (int p = 0; p < PX; ++p) { accN2[o][p] = 0ull; accD2[o][p] = 0ull; }

#pragma unroll 1
    for (int ch = 0; ch < NCHUNK; ++ch, ++q) {
      const int cur = q & 1;
      cp_async_wait_all();
      fixup(ch, cur);
      __syncthreads();                        // chunk q visible; everyone is done with buffer cur^1 (chunk q-1)
      if (ch + 1 < NCHUNK) issue(t, ch + 1, cur ^ 1);
      else if (t + (int)gridDim.x < ntiles) issue(t + gridDim.x, 0, cur ^ 1);

      const float* s_xc = s_buf + (size_t)cur * 2 * SM::TILE_FLOATS;
      const float* s_c = s_xc + SM::TILE_FLOATS;
#pragma unroll 1
      for (int cc = 0; cc < CCH; ++cc) {
        const float* wrow = s_w + (size_t)(ch * CCH + cc) * K * K * CO;
#pragma unroll
        for (int u = 0; u < K; ++u) {
          float xr[4 * NV], cr[4 * NV];
          const int so = (cc * RH + ty + u) * SW + 4 * tx;
#pragma unroll
          for (int qv = 0; qv < NV; ++qv) {
            VecLoad<4>::ld(xr + 4 * qv, s_xc + so + 4 * qv);
            VecLoad<4>::ld(cr + 4 * qv, s_c + so + 4 * qv);
          }
#pragma unroll
          for (int v = 0; v < K; ++v) {
            u64 w2[CO_P];
            if constexpr (CO % 4 == 0) {
#pragma unroll
              for (int o = 0; o < CO_P; o += 2) {
                const ulonglong2 tw = *reinterpret_cast<const ulonglong2*>(wrow + (u * K + v) * CO + 2 * o);
                w2[o] = tw.x; w2[o + 1] = tw.y;
              }
            } else {
#pragma unroll
              for (int o = 0; o < CO_P; ++o) w2[o] = *reinterpret_cast<const u64*>(wrow + (u * K + v) * CO + 2 * o);
            }
#pragma unroll
            for (int o = 0; o < CO_P; ++o)
#pragma unroll
              for (int p = 0; p < PX; ++p) {
                ffma2(accN2[o][p], w2[o], xr[OFF + p + v]);
                ffma2(accD2[o][p], w2[o], cr[OFF + p + v]);
              }
          }
        }
      }
    }

    // ---- epilogue of tile t (the next tile's first chunk is already in flight)
    float aN[CO][PX], aD[CO][PX];
#pragma unroll
    for (int o = 0; o < CO; ++o)
#pragma unroll
      for (int p = 0; p < PX; ++p) {
        aN[o][p] = (o & 1) ? hi32f(accN2[o >> 1][p]) : lo32f(accN2[o >> 1][p]);
        aD[o][p] = (o & 1) ? hi32f(accD2[o >> 1][p]) : lo32f(accD2[o >> 1][p]);
      }
    const int oy = oy0 + ty, ox = ox0 + PX * tx;
    if constexpr (TAIL) {
      // nconv6 (valid 3x3 evaluated centred) -> nconv7 (1x1, pad-2 ring = bias7) -> crop: step1.py:90-94
      if (oy < a.H) {
        const float b7 = __ldg(a.b7);
        float res[PX];
#pragma unroll
        for (int p = 0; p < PX; ++p) {
          float n7 = 0.f, d7 = 0.f;
#pragma unroll
          for (int o = 0; o < CO; ++o) {
            const float D = aD[o][p];
            const float x6 = __fdividef(aN[o][p], D + a.eps) + s_bias[o];
            const float c6 = D * s_sum[o];
            const float w7 = __ldg(a.w7 + o);
            n7 = fmaf(w7, x6 * c6, n7);
            d7 = fmaf(w7, c6, d7);
          }
          const int gx = ox + p;
          const bool ring = (oy == 0) || (oy == a.H - 1) || (gx == 0) || (gx >= a.W - 1);
          res[p] = ring ? b7 : (__fdividef(n7, d7 + a.eps) + b7);
        }
        if (ox + 3 < a.W)
          *reinterpret_cast<float4*>(a.out + ((size_t)b * a.H + oy) * a.W + ox) = make_float4(res[0], res[1], res[2], res[3]);
      }
    } else {
      float yv[CO][PX], cv[CO][PX];
#pragma unroll
      for (int o = 0; o < CO; ++o) {
        const float bo = s_bias[o], so = s_sum[o];
#pragma unroll
        for (int p = 0; p < PX; ++p) {
          const float D = aD[o][p];
          yv[o][p] = __fdividef(aN[o][p], D + a.eps) + bo;
          cv[o][p] = D * so;
        }
      }
      if (oy < a.HO && ox + 3 < a.WO) {
#pragma unroll
        for (int o = 0; o < CO; ++o) {
          const size_t base = (((size_t)b * CO_ALL + cog * CO + o) * a.HO + oy) * a.WO + ox;
          *reinterpret_cast<float4*>(a.y + base) = make_float4(yv[o][0], yv[o][1], yv[o][2], yv[o][3]);
          *reinterpret_cast<float4*>(a.co + base) = make_float4(cv[o][0], cv[o][1], cv[o][2], cv[o][3]);
        }
      }
      if constexpr (POOL) {
        // independent 2x2/2 max-pools of y and cout (step1.py:62-74); the lower row of a window lives in lane^16
        const int py = oy >> 1;
#pragma unroll
        for (int o = 0; o < CO; ++o) {
          float ylo[PX], clo[PX];
#pragma unroll
          for (int p = 0; p < PX; ++p) {
            ylo[p] = __shfl_xor_sync(0xffffffffu, yv[o][p], 16);
            clo[p] = __shfl_xor_sync(0xffffffffu, cv[o][p], 16);
          }
          if ((ty & 1) == 0 && py < a.HP) {
            const size_t base = (((size_t)b * CO_ALL + cog * CO + o) * a.HP + py) * a.WP;
#pragma unroll
            for (int h = 0; h < PX / 2; ++h) {
              const int pxo = (ox >> 1) + h;
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
  }
  cp_async_wait_all();
}

// Eligibility (checked by the caller): forward, a.vec_ok, thr >= 0 under F_THRESH, source B an exact 2x upsample
// with 8-byte rows (a.up2_vec) when CB > 0.
template <int CA, int CB, int CO_ALL, int CO, int K, int PAD, int FLAGS>
static int launch_pipe(const TiledArgs& a, cudaStream_t st) {
  using SM = PipeSmem<CA, CB, CO, K>;
  constexpr int NG = CO_ALL / CO;
  auto kern = nconv_pipe<CA, CB, CO_ALL, CO, K, PAD, FLAGS>;
  static bool attr_done[NCONV_MAX_DEVICES] = {};
  static int sm_count[NCONV_MAX_DEVICES] = {};
  int dev = 0;
  NCONV_RETURN_IF_ERR(cudaGetDevice(&dev));
  if (dev >= NCONV_MAX_DEVICES) return NCONV_ERR_UNSUPPORTED;
  if (!attr_done[dev]) {
    NCONV_RETURN_IF_ERR(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SM::BYTES));
    NCONV_RETURN_IF_ERR(cudaDeviceGetAttribute(&sm_count[dev], cudaDevAttrMultiProcessorCount, dev));
    attr_done[dev] = true;
  }
  constexpr bool TAIL = (FLAGS & NCONV_B200_F_TAIL) != 0;
  constexpr bool THRESH = (FLAGS & NCONV_B200_F_THRESH) != 0;
  constexpr bool POOL = (FLAGS & NCONV_B200_F_POOL) != 0;
  const int ow = TAIL ? a.W : a.WO, oh = TAIL ? a.H : a.HO;
  const int tiles_x = ceil_div(ow, TW), tiles_y = ceil_div(oh, TH);
  const long long ntiles = (long long)tiles_x * tiles_y * a.B;
  if (ntiles > 0x7fffffffLL) return NCONV_ERR_UNSUPPORTED;
  const long long slots = 2LL * sm_count[dev] / NG > 0 ? 2LL * sm_count[dev] / NG : 1;
  const dim3 grid((unsigned)(ntiles < slots ? ntiles : slots), NG);
  const double px_in = (double)a.B * a.H * a.W, px_out = (double)a.B * oh * ow;
  double bytes, flops;
  if (TAIL) {
    bytes = 4.0 * (2.0 * CA * px_in + 2.0 * CB * a.B * a.HB * a.WB + px_out);
    flops = 4.0 * (CA + CB) * CO_ALL * K * K * px_out + 4.0 * CO_ALL * px_out;
  } else {
    bytes = 4.0 * ((THRESH ? 1.0 : 2.0) * CA * px_in + 2.0 * CB * a.B * a.HB * a.WB + 2.0 * CO_ALL * px_out +
                   (POOL ? 2.0 * CO_ALL * a.B * a.HP * a.WP : 0.0));
    flops = 4.0 * (CA + CB) * CO_ALL * K * K * px_out;
  }
  char label[64];
  snprintf(label, sizeof(label), "nconv_fwd %d+%d->%d k%d f%d", CA, CB, CO_ALL, K, FLAGS & 7);
  {
    LaunchScope ls(st, label, flops, bytes);
    kern<<<grid, NTHREADS, SM::BYTES, st>>>(a, tiles_x, tiles_y, (int)ntiles);
  }
  return (int)cudaGetLastError();
}

}  // namespace nconv
