// Tensor-core NConv2d for DNET's FMA-bound layers (/root/reference/models/step1.py:38-49: the four 8->8 5x5 layers
// and the three [8|8]->8 3x3 layers; forward = step1.py:116-149) on tcgen05, plus the small CUDA-core kernels around
// the N8 layout (n8.cuh): ring fill, pack / unpack, and nconv1 (1->8 5x5 with the threshold prologue).
//
// Why pixels on M with the horizontal taps on N.  On CUDA cores these layers are FMA-bound (8->8 5x5: 3 200 FMA per
// pixel against 128 B).  As an implicit GEMM the contraction is tiny per output (8 channels), so the
// MMA must produce more than 8 useful columns per issue:
//     D[128 pixels][N = K*8] += A[128 pixels][16] * B[N][16]^T
//   * A = 128 consecutive pixels of a staged N8 tile (flat, pitch 32 pixels: one staged row per epilogue warp);
//   * B row n = dx*8 + co holds W[co][ci][dy][dx] for the MMA's vertical tap dy: the K (=5 or 3) horizontal taps are
//     stacked on N, so column (dx, co) of pixel m is tap dx's contribution to OUTPUT pixel m - dx ... i.e. the
//     epilogue adds K lane-shifted columns (warp shuffles: a warp owns one staged row, 32 - (K-1) outputs);
//   * the vertical taps and the three products of the split-bf16 (bf16x3) scheme are stacked on K, two 8-channel
//     chunks per MMA: [x_hi@dy ; x_lo@dy] against [W_hi[dy] ; W_hi[dy]] (one MMA per dy), and the W_lo * x_hi terms
//     paired over dy (5x5) or over the two sources (3x3).  8 (5x5) / 9 (3x3, two sources) MMAs per stream per 128 pixels.
//   * both streams (data*conf -> N, conf -> D) use the same B tiles and separate accumulator columns.
// The accumulators live in a TMEM ring of 5 (5x5: 96 columns) or 8 (3x3: 64 columns) M-tile slots; four epilogue warps
// drain a slot (tcgen05.ld), do the shift-sum, N/(D+eps)+b and D/sum(W) (step1.py:123-147), and write whichever
// outputs the layer needs: N8 at the same size, the 2x2 max-pooled N8 (step1.py:62-74), the 2x nearest-upsampled N8
// (step1.py:78-89), fp32 NCHW, or the nconv7 + crop tail (step1.py:92-94).
#include "common.cuh"
#include "launch_count.h"
#include "n8.cuh"
#include "nconv_tc.h"
#include "nconv_b200.h"
#include "p8.cuh"
#include "tc_ptx.cuh"
#include "tc_tmap.cuh"
#include <cuda.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace nconv {

// ---------------------------------------------------------------------------------------------------------------
// N8 helpers (CUDA cores, elementwise)
// ---------------------------------------------------------------------------------------------------------------
__global__ void n8_zero_ring_kernel(__nv_bfloat16* buf, N8Geom g) {
  // one thread per (plane, ring pixel): 2 top + 2 bottom rows, 2 left + 2 right columns
  const int ring_px = 2 * N8_RING * g.Wp + 2 * N8_RING * g.H;
  const long long total = (long long)g.B * 4 * ring_px;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
    const int r = (int)(idx % ring_px);
    const long long pl = idx / ring_px;
    int y, x;
    if (r < 2 * N8_RING * g.Wp) {
      const int rr = r / g.Wp; x = r % g.Wp;
      y = rr < N8_RING ? rr : g.Hp - 2 * N8_RING + rr;
    } else {
      const int q = r - 2 * N8_RING * g.Wp;
      const int cc = q % (2 * N8_RING);
      y = N8_RING + q / (2 * N8_RING);
      x = cc < N8_RING ? cc : g.Wp - 2 * N8_RING + cc;
    }
    *reinterpret_cast<uint4*>(buf + pl * g.plane_elems() + ((size_t)y * g.Wp + x) * 8) = make_uint4(0, 0, 0, 0);
  }
}

// thread = pixel: 8 channels of x and c -> xc = x*c, c, each split hi/lo
__global__ void n8_pack_kernel(const float* __restrict__ x, const float* __restrict__ c, __nv_bfloat16* __restrict__ out, N8Geom g) {
  const long long total = (long long)g.B * g.H * g.W;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
    const int px = (int)(idx % g.W);
    const int py = (int)((idx / g.W) % g.H);
    const int n = (int)(idx / ((long long)g.W * g.H));
    float xc[8], cc[8];
#pragma unroll
    for (int ch = 0; ch < 8; ++ch) {
      const size_t o = (((size_t)n * 8 + ch) * g.H + py) * g.W + px;
      cc[ch] = __ldg(c + o);
      xc[ch] = __ldg(x + o) * cc[ch];
    }
    uint4 h, l;
    const size_t po = g.pixel(py, px);
    split8(xc, h, l);
    *reinterpret_cast<uint4*>(out + g.plane(n, 0, 0) + po) = h;
    *reinterpret_cast<uint4*>(out + g.plane(n, 0, 1) + po) = l;
    split8(cc, h, l);
    *reinterpret_cast<uint4*>(out + g.plane(n, 1, 0) + po) = h;
    *reinterpret_cast<uint4*>(out + g.plane(n, 1, 1) + po) = l;
  }
}

__global__ void n8_unpack_kernel(const __nv_bfloat16* __restrict__ in, float* __restrict__ xc, float* __restrict__ c, N8Geom g) {
  const long long total = (long long)g.B * g.H * g.W;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
    const int px = (int)(idx % g.W);
    const int py = (int)((idx / g.W) % g.H);
    const int n = (int)(idx / ((long long)g.W * g.H));
    const size_t po = g.pixel(py, px);
    float a[8], b[8];
    join8(*reinterpret_cast<const uint4*>(in + g.plane(n, 0, 0) + po), *reinterpret_cast<const uint4*>(in + g.plane(n, 0, 1) + po), a);
    join8(*reinterpret_cast<const uint4*>(in + g.plane(n, 1, 0) + po), *reinterpret_cast<const uint4*>(in + g.plane(n, 1, 1) + po), b);
#pragma unroll
    for (int ch = 0; ch < 8; ++ch) {
      const size_t o = (((size_t)n * 8 + ch) * g.H + py) * g.W + px;
      if (xc) xc[o] = a[ch];
      if (c) c[o] = b[ch];
    }
  }
}

// ---------------------------------------------------------------------------------------------------------------
// nconv1 (step1.py:53-56): c0 = (S > thr); NConv2d 1 -> 8, 5x5, pad 2; result written as N8.
// HBM-bound by construction (4 B in, 64 B out per pixel).  A 256-thread CTA owns a 32 x 8 pixel tile: the raw (12 x 36)
// halo tile of S is staged once (zero outside the image: conf = 0 there, exactly the zero padding), a thread computes
// all 8 output channels of ONE pixel with FFMA2 over channel pairs (weights broadcast from shared memory) and writes
// its four 16-byte N8 vectors (consecutive lanes = consecutive pixels: 512-byte runs).
// ---------------------------------------------------------------------------------------------------------------
constexpr int N1_TW = 32, N1_TH = 16, N1_PY = 2;     // a thread owns rows ty and ty+8 of its column: weight reads shared by 2 pixels

__global__ void __launch_bounds__(256) nconv1_n8_kernel(const float* __restrict__ S, const float* __restrict__ w, const float* __restrict__ bias,
                                                        __nv_bfloat16* __restrict__ out, N8Geom g, float thr, float eps) {
  __shared__ float2 s_in[N1_TH + 4][N1_TW + 4];                   // (S*conf, conf) with conf = S > thr; zero outside the image
  __shared__ __align__(16) float s_w[25][8];
  __shared__ float s_rs[8], s_b[8];
  const int tid = threadIdx.x;
  const int tx = tid & 31, ty = tid >> 5;
  const int x0 = blockIdx.x * N1_TW, y0 = blockIdx.y * N1_TH, n = blockIdx.z;
  if (tid < 200) s_w[tid % 25][tid / 25] = __ldg(w + tid);          // w [8][1][5][5] -> [tap][co]
  for (int i = tid; i < (N1_TH + 4) * (N1_TW + 4); i += 256) {
    const int r = i / (N1_TW + 4), cc = i % (N1_TW + 4);
    const int gy = y0 - 2 + r, gx = x0 - 2 + cc;
    const float x = (gy >= 0 && gy < g.H && gx >= 0 && gx < g.W) ? __ldg(S + ((size_t)n * g.H + gy) * g.W + gx) : 0.f;
    const float c = x > thr ? 1.f : 0.f;
    s_in[r][cc] = make_float2(x * c, c);
  }
  __syncthreads();
  if (tid < 8) {
    float s = 0.f;
    for (int t = 0; t < 25; ++t) s += s_w[t][tid];
    s_rs[tid] = 1.f / s;
    s_b[tid] = __ldg(bias + tid);
  }
  u64 aN[N1_PY][4], aD[N1_PY][4];
#pragma unroll
  for (int p = 0; p < N1_PY; ++p)
#pragma unroll
    for (int o = 0; o < 4; ++o) { aN[p][o] = 0ull; aD[p][o] = 0ull; }
#pragma unroll
  for (int u = 0; u < 5; ++u)
#pragma unroll
    for (int v = 0; v < 5; ++v) {
      const ulonglong2 w01 = *reinterpret_cast<const ulonglong2*>(&s_w[u * 5 + v][0]);
      const ulonglong2 w23 = *reinterpret_cast<const ulonglong2*>(&s_w[u * 5 + v][4]);
#pragma unroll
      for (int p = 0; p < N1_PY; ++p) {
        const float2 xc = s_in[ty + 8 * p + u][tx + v];
        ffma2(aN[p][0], w01.x, xc.x); ffma2(aN[p][1], w01.y, xc.x); ffma2(aN[p][2], w23.x, xc.x); ffma2(aN[p][3], w23.y, xc.x);
        ffma2(aD[p][0], w01.x, xc.y); ffma2(aD[p][1], w01.y, xc.y); ffma2(aD[p][2], w23.x, xc.y); ffma2(aD[p][3], w23.y, xc.y);
      }
    }
  __syncthreads();
  const int ox = x0 + tx;
  if (ox >= g.W) return;
#pragma unroll
  for (int p = 0; p < N1_PY; ++p) {
    const int oy = y0 + ty + 8 * p;
    if (oy >= g.H) continue;
    float xc[8], cc[8];
#pragma unroll
    for (int o = 0; o < 8; ++o) {
      const float N = (o & 1) ? hi32f(aN[p][o >> 1]) : lo32f(aN[p][o >> 1]);
      const float D = (o & 1) ? hi32f(aD[p][o >> 1]) : lo32f(aD[p][o >> 1]);
      const float y = __fdividef(N, D + eps) + s_b[o];
      cc[o] = D * s_rs[o];
      xc[o] = y * cc[o];
    }
    uint4 h, l;
    const size_t po = g.pixel(oy, ox);
    split8_fast(xc, h, l);
    *reinterpret_cast<uint4*>(out + g.plane(n, 0, 0) + po) = h;
    *reinterpret_cast<uint4*>(out + g.plane(n, 0, 1) + po) = l;
    split8_fast(cc, h, l);
    *reinterpret_cast<uint4*>(out + g.plane(n, 1, 0) + po) = h;
    *reinterpret_cast<uint4*>(out + g.plane(n, 1, 1) + po) = l;
  }
}

}  // namespace nconv

namespace nconv {

// ---------------------------------------------------------------------------------------------------------------
// tcgen05 NConv2d
// ---------------------------------------------------------------------------------------------------------------
constexpr int NT_ESETS = 3;                 // epilogue sets of 4 warps; set e drains the M-tiles with (running index) % NT_ESETS == e
constexpr int NT_ISSUERS = 2;               // MMA issuer warps (1, 2): issuer i takes the M-tiles with (running index) % NT_ISSUERS == i
constexpr int NT_THREADS = 128 + 128 * NT_ESETS;  // warp 0: producer, warps 1-2: MMA issuers, warp 3: idle, then the epilogue warps (TMEM lane quarter = warp % 4)
constexpr int NT_PX_BYTES = NT_ESETS * 2 * 4 * 8 * 32 * 4;    // pool exchange: [set][parity][staged row in M-tile][4 y | 4 c][lane] floats
constexpr int NT_P = 32;                    // staged row pitch in pixels: one staged row per epilogue warp
constexpr int NT_ROW_BYTES = NT_P * 16;     // one staged row of one plane
constexpr int NT_SMEM_MAX = 227 * 1024 - 4 * 1024;    // dynamic part (static: barriers and per-channel constants)
constexpr int NT_SMEM_MIN = 120 * 1024;     // > half an SM's shared memory: one CTA per SM (each allocates all 512 TMEM columns)

template <int K> struct NtCfg {
  static constexpr int PAD = K / 2;
  static constexpr int HALO = K - 1;
  static constexpr int WT = NT_P - HALO;              // output pixels per staged row
  static constexpr int NUSE = K * 8;                  // useful B rows: (dx, co)
  static constexpr int NROW = (NUSE + 15) / 16 * 16;  // M = 128 needs N % 16 == 0
  static constexpr int SLOT_COLS = 2 * NROW;          // N-stream | D-stream accumulators of one M-tile
  static constexpr int NSLOT = 512 / SLOT_COLS;       // TMEM ring depth: 5 (5x5) / 8 (3x3)
  static constexpr int WBLK = NROW * 16;              // bytes of one weight block [NROW][8] bf16
};

struct NtArgs {
  const __nv_bfloat16* wpack; const float* w; const float* bias;
  __nv_bfloat16* out_n8; __nv_bfloat16* out_pool; __nv_bfloat16* out_up; float* y; float* cout;
  const float* w_tail; const float* b_tail; float* out_tail;
  int B, H, W, CI;
  int TM, rows_ld, tiles_x, tiles_y, ntiles;
  int issuers, esets;          // active MMA issuer warps (<= NT_ISSUERS) / epilogue sets (<= NT_ESETS)
  int dbg;                     // NCONV_B200_NT_DEBUG bits (timing experiments only): 1 no global stores, 2 no pool, 4 no MMAs, 8 no shift-sum
  float eps;
};

__device__ __forceinline__ void tmem_ld8(uint32_t taddr, uint32_t* r) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]) : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t* r) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                 "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]) : "r"(taddr));
}
template <int K> __device__ __forceinline__ void nt_load_stream(uint32_t taddr, uint32_t* v) {
  if constexpr (K == 5) { tmem_ld32(taddr, v); tmem_ld8(taddr + 32, v + 32); }
  else { tmem_ld16(taddr, v); tmem_ld8(taddr + 16, v + 16); }
}

// (data*conf, conf) of one pixel as its four 16-byte N8 vectors
struct N8Px { uint4 xh, xl, ch, cl; };
__device__ __forceinline__ N8Px n8_make(const float* xc, const float* cc) {
  N8Px p;
  split8_fast(xc, p.xh, p.xl);
  split8_fast(cc, p.ch, p.cl);
  return p;
}
__device__ __forceinline__ void n8_put(__nv_bfloat16* buf, const N8Geom& g, int n, size_t po, const N8Px& p) {
  __nv_bfloat16* b0 = buf + g.plane(n, 0, 0) + po;
  const size_t pe = g.plane_elems();
  *reinterpret_cast<uint4*>(b0) = p.xh;
  *reinterpret_cast<uint4*>(b0 + pe) = p.xl;
  *reinterpret_cast<uint4*>(b0 + 2 * pe) = p.ch;
  *reinterpret_cast<uint4*>(b0 + 3 * pe) = p.cl;
}

// Debug aid: nconv_b200_debug_buffer() hands the library 64 bytes of pinned host memory; a wait that times out records
// {code, block, warp, extra} there before trapping, so the host can tell which barrier of which role gave up.
__device__ unsigned int* g_nt_dbg = nullptr;
// The loop exit is a warp vote, i.e. WARP-UNIFORM: every lane leaves the wait in the same iteration.  This matters for
// the MMA issuer warps, whose code after the wait is warp-convergent and runs on the uniform datapath (one UTCHMMA /
// commit per warp execution): with a per-lane exit (`while (!try_wait)`) the compiler emits no reconvergence point, the
// lanes that saw the phase flip first ran ahead as their own group and the MMAs and commits were issued once per group
// - duplicate barrier arrivals, and a deadlock that showed up only in long runs (B >= 34 at 352x1216).
__device__ __forceinline__ void nt_wait(uint32_t bar, uint32_t parity, unsigned code, unsigned extra) {
  uint32_t spins = 0;
  while (!__all_sync(0xffffffffu, mbar_try_wait(bar, parity))) {
    if (++spins > (1u << 22)) {
      unsigned int* d = g_nt_dbg;
      if (d && atomicCAS(d, 0u, code) == 0u) {
        d[1] = blockIdx.x; d[2] = threadIdx.x >> 5; d[3] = extra; d[4] = parity;
        __threadfence_system();
      }
      __trap();
    }
  }
}

template <int K, int NSRC>
__global__ void __launch_bounds__(NT_THREADS, 1) nconv_tc_kernel(const __grid_constant__ NtArgs a, const __grid_constant__ CUtensorMap tm0,
                                                                 const __grid_constant__ CUtensorMap tm1) {
  using C = NtCfg<K>;
  constexpr int NBLK = 2 * NSRC * K + 1;                       // hi/lo block per (source, dy) + one zero block
  constexpr uint32_t W_BYTES = (NBLK * C::WBLK + 127) / 128 * 128;
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t s_full[2], s_empty[2], t_full[C::NSLOT], t_empty[C::NSLOT], s_wfull;
  __shared__ uint32_t s_tmem_base;
  __shared__ float s_rs[8], s_b[8], s_w7[8];

  // the shuffle makes the warp index provably warp-uniform, so role branches are uniform branches and the MMA issue
  // code below compiles to the uniform datapath
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
  const uint32_t plane_bytes = (uint32_t)a.rows_ld * NT_ROW_BYTES;
  const uint32_t stage_bytes = NSRC * 4 * plane_bytes;
  const uint32_t smem_base = smem_u32(smem);
  const uint32_t ring = smem_base + W_BYTES + NT_PX_BYTES;
  float* s_px = reinterpret_cast<float*>(smem + W_BYTES);

  if (threadIdx.x < 8) {
    // cout = D / sum(W[o]) (step1.py:141-147) over the ORIGINAL fp32 weights
    float s = 0.f;
    const int nw = a.CI * K * K;
    for (int j = 0; j < nw; ++j) s += __ldg(a.w + (size_t)threadIdx.x * nw + j);
    s_rs[threadIdx.x] = 1.f / s;
    s_b[threadIdx.x] = __ldg(a.bias + threadIdx.x);
    s_w7[threadIdx.x] = a.w_tail ? __ldg(a.w_tail + threadIdx.x) : 0.f;
  }
  if (threadIdx.x == 0) {
    for (int s = 0; s < 2; ++s) { mbar_init(smem_u32(&s_full[s]), 1); mbar_init(smem_u32(&s_empty[s]), a.issuers); }
    for (int s = 0; s < C::NSLOT; ++s) { mbar_init(smem_u32(&t_full[s]), 1); mbar_init(smem_u32(&t_empty[s]), 4); }
    mbar_init(smem_u32(&s_wfull), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem_base)), "r"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = s_tmem_base;
  const int rows_tile = 4 * a.TM;

  if (warp == 0) {
    // ===================== producer =====================
    if (lane == 0) {
      const uint32_t wf = smem_u32(&s_wfull);
      mbar_arrive_expect_tx(wf, NBLK * C::WBLK);
      bulk_g2s(smem_base, a.wpack, NBLK * C::WBLK, wf);
    }
    uint32_t stage = 0, phase = 0;
    for (int tile = blockIdx.x; tile < a.ntiles; tile += gridDim.x) {
      const int tx = tile % a.tiles_x, ty = (tile / a.tiles_x) % a.tiles_y, n = tile / (a.tiles_x * a.tiles_y);
      const int y0 = ty * rows_tile, x0 = tx * C::WT;
      nt_wait(smem_u32(&s_empty[stage]), phase ^ 1, 1, tile);
      const uint32_t full = smem_u32(&s_full[stage]);
      if (lane == 0) mbar_arrive_expect_tx(full, stage_bytes);
      __syncwarp();
      if (lane < NSRC * 4) {
        // one tensor copy per staged plane: box [32 px as 128 words][rows_ld rows] of plane n*4 + (stream*2 + part);
        // staged (0, 0) is image (y0 - PAD, x0 - PAD) = padded (y0 - PAD + 2, x0 - PAD + 2); out-of-range parts are zero-filled
        const int src = lane >> 2, pl = lane & 3;
        const CUtensorMap* tm = src == 0 ? &tm0 : &tm1;
        asm volatile(
            "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
            ::"r"(ring + stage * stage_bytes + lane * plane_bytes), "l"(reinterpret_cast<uint64_t>(tm)),
              "r"(4 * (x0 - C::PAD + N8_RING)), "r"(y0 - C::PAD + N8_RING), "r"(n * 4 + pl), "r"(full)
            : "memory");
      }
      __syncwarp();
      stage ^= 1; if (stage == 0) phase ^= 1;
    }
  } else if (warp == 1 || (warp == 2 && a.issuers == 2)) {
    // ===================== MMA issuers =====================
    // Issuing one MMA costs ~100 cycles of the issuing warp (descriptor moves to uniform registers, elect, branch), more
    // than the MMA itself at N = 32..48, so two warps issue for alternate M-tiles (independent accumulator slots).
    const int issuer = warp - 1;
    constexpr uint32_t IDESC = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(C::NROW >> 3) << 17) | ((128u >> 4) << 24);
    constexpr uint32_t DESC_HI = (128u >> 4) | (1u << 14);      // SBO = 128 B, descriptor version 1
    nt_wait(smem_u32(&s_wfull), 0, 2, 0);
    tc_fence_after();
    uint32_t stage = 0, phase = 0, it = 0, slot = 0, sphase = 0;
    const uint32_t w16 = smem_base >> 4;
    const uint32_t plane16 = plane_bytes >> 4;
    const uint32_t lbo_hl = plane16 << 16;                      // hi plane -> lo plane of the same stream
    constexpr uint32_t ROW16 = NT_ROW_BYTES >> 4, WBLK16 = C::WBLK >> 4;
    // The whole issue loop is warp-convergent (every lane runs it on warp-uniform values, one lane is elected inside
    // umma_bf16_ws): descriptors stay in uniform registers, UTCHMMAs issue back to back.
    for (int tile = blockIdx.x; tile < a.ntiles; tile += gridDim.x) {
      const int ty = (tile / a.tiles_x) % a.tiles_y;
      const int rows_left = a.H - ty * rows_tile;
      const int tmv = rows_left >= rows_tile ? a.TM : (rows_left + 3) / 4;       // M-tiles holding image rows
      nt_wait(smem_u32(&s_full[stage]), phase, 3, tile);
      tc_fence_after();
      const uint32_t st16 = (ring + stage * stage_bytes) >> 4;
      for (int t = 0; t < tmv; ++t) {
        if ((int)(it & (uint32_t)(a.issuers - 1)) == issuer) {       // issuers is 1 or 2
          nt_wait(smem_u32(&t_empty[slot]), sphase ^ 1, 4, it);
          tc_fence_after();
          if (!(a.dbg & 4))
#pragma unroll
          for (int s = 0; s < 2; ++s) {
            const uint32_t d = tmem_base + slot * C::SLOT_COLS + s * C::NROW;
            const uint32_t xs = st16 + ((uint32_t)(s * 2) * a.rows_ld + 4 * t) * ROW16;     // stream s, part hi, first row of the M-tile
            // [x_hi@dy ; x_lo@dy] x [W_hi[dy] ; W_hi[dy]]  (B's two K-halves alias one block: LBO = 0)
#pragma unroll
            for (int src = 0; src < NSRC; ++src)
#pragma unroll
              for (int dy = 0; dy < K; ++dy) {
                const uint32_t ax = xs + (uint32_t)src * 4 * plane16 + dy * ROW16;
                const uint32_t bw = w16 + ((src * K + dy) * 2) * WBLK16;
                umma_bf16_ws(d, ((uint64_t)DESC_HI << 32) | (ax | lbo_hl), ((uint64_t)DESC_HI << 32) | bw, IDESC,
                             (src == 0 && dy == 0) ? 0u : 1u);
              }
            // x_hi x W_lo, two chunks per MMA
            if constexpr (NSRC == 1) {
#pragma unroll
              for (int dy = 0; dy < K; dy += 2) {
                constexpr uint32_t A_LBO_PAIR = ROW16 << 16;
                const bool pair = dy + 1 < K;
                const uint32_t ax = xs + dy * ROW16;
                const uint32_t bw = w16 + (dy * 2 + 1) * WBLK16;
                // paired: second K-half = next staged row against W_lo[dy+1]; last odd tap: both halves read row dy,
                // the second against the zero block (which follows W_lo[K-1] in the packed weights)
                const uint32_t a_lbo = pair ? A_LBO_PAIR : 0u;
                const uint32_t b_lbo = (uint32_t)((pair ? 2 : 1) * WBLK16) << 16;
                umma_bf16_ws(d, ((uint64_t)DESC_HI << 32) | (ax | a_lbo), ((uint64_t)DESC_HI << 32) | (bw | b_lbo), IDESC, 1u);
              }
            } else {
#pragma unroll
              for (int dy = 0; dy < K; ++dy) {
                const uint32_t ax = xs + dy * ROW16;
                const uint32_t bw = w16 + (dy * 2 + 1) * WBLK16;
                const uint32_t a_lbo = (4 * plane16) << 16;                            // same plane of source B
                constexpr uint32_t B_LBO = (uint32_t)(2 * K * WBLK16) << 16;           // W_lo[B][dy]
                umma_bf16_ws(d, ((uint64_t)DESC_HI << 32) | (ax | a_lbo), ((uint64_t)DESC_HI << 32) | (bw | B_LBO), IDESC, 1u);
              }
            }
          }
          umma_commit_ws(smem_u32(&t_full[slot]));
        }
        ++it;
        if (++slot == (uint32_t)C::NSLOT) { slot = 0; sphase ^= 1; }
      }
      umma_commit_ws(smem_u32(&s_empty[stage]));                 // every MMA of this issuer reading the stage has retired
      stage ^= 1; if (stage == 0) phase ^= 1;
    }
  } else if (warp >= 4 && warp < 4 + 4 * a.esets) {
    // ===================== epilogue (warps 4..) =====================
    // One warp per SM sub-partition cannot hide its own instruction latencies (ncu/events: 2 090 cycles per M-tile with a
    // single set against 512 of MMAs), so NT_ESETS sets of four warps work on alternate M-tiles.
    const int q = warp & 3;                                      // TMEM lane quarter = staged row of the M-tile
    const int set = (warp - 4) >> 2;
    const N8Geom gs = N8Geom::make(a.B, a.H, a.W);
    const N8Geom gp = N8Geom::make(a.B, a.H / 2, a.W / 2);
    const N8Geom gu = N8Geom::make(a.B, 2 * a.H, 2 * a.W);
    const float b7 = a.b_tail ? __ldg(a.b_tail) : 0.f;
    uint32_t it = 0;
    int erot = 0;
    for (int tile = blockIdx.x; tile < a.ntiles; tile += gridDim.x) {
      const int tx = tile % a.tiles_x, ty = (tile / a.tiles_x) % a.tiles_y, n = tile / (a.tiles_x * a.tiles_y);
      const int y0 = ty * rows_tile, x0 = tx * C::WT;
      const int rows_left = a.H - y0;
      const int tmv = rows_left >= rows_tile ? a.TM : (rows_left + 3) / 4;
      for (int t = 0; t < tmv; ++t, ++it) {
        const bool mine = (erot == set);                         // it % esets == set, without the integer division
        if (++erot == a.esets) erot = 0;
        if (!mine) continue;
        const uint32_t slot = it % C::NSLOT, sphase = (it / C::NSLOT) & 1;
        nt_wait(smem_u32(&t_full[slot]), sphase, 5, it);
        tc_fence_after();
        const uint32_t tcol = tmem_base + ((uint32_t)(q * 32) << 16) + slot * C::SLOT_COLS;
        uint32_t vN[C::NUSE], vD[C::NUSE];
        nt_load_stream<K>(tcol, vN);
        nt_load_stream<K>(tcol + C::NROW, vD);
        tmem_ld_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_u32(&t_empty[slot]));    // the slot is in registers: the next M-tile may overwrite it
        // column (dx, co) of staged pixel l is tap dx's term of output pixel l - dx: add K lane-shifted columns
        float N[8], D[8];
#pragma unroll
        for (int co = 0; co < 8; ++co) { N[co] = __uint_as_float(vN[co]); D[co] = __uint_as_float(vD[co]); }
        if (!(a.dbg & 8))
#pragma unroll
        for (int dx = 1; dx < K; ++dx)
#pragma unroll
          for (int co = 0; co < 8; ++co) {
            N[co] += __shfl_down_sync(0xffffffffu, __uint_as_float(vN[dx * 8 + co]), dx);
            D[co] += __shfl_down_sync(0xffffffffu, __uint_as_float(vD[dx * 8 + co]), dx);
          }
        const int oy = y0 + 4 * t + q, ox = x0 + lane;
        const bool valid = lane < C::WT && oy < a.H && ox < a.W && !(a.dbg & 1);
        float yv[8], cv[8];
#pragma unroll
        for (int co = 0; co < 8; ++co) {
          yv[co] = __fdividef(N[co], D[co] + a.eps) + s_b[co];   // step1.py:123-131
          cv[co] = D[co] * s_rs[co];                             // step1.py:141-147
        }
        if (a.out_tail) {
          // nconv7 (1x1 over the 8 channels; its pad-2 ring of zeros gives exactly bias7 on the border) + crop, step1.py:92-94
          float n7 = 0.f, d7 = 0.f;
#pragma unroll
          for (int co = 0; co < 8; ++co) { n7 = fmaf(s_w7[co], yv[co] * cv[co], n7); d7 = fmaf(s_w7[co], cv[co], d7); }
          const bool ringpx = (oy == 0) || (oy == a.H - 1) || (ox == 0) || (ox >= a.W - 1);
          if (valid) a.out_tail[((size_t)n * a.H + oy) * a.W + ox] = ringpx ? b7 : (__fdividef(n7, d7 + a.eps) + b7);
        }
        float xc[8];
#pragma unroll
        for (int co = 0; co < 8; ++co) xc[co] = yv[co] * cv[co];
        if (valid) {
          if (a.y) {
#pragma unroll
            for (int co = 0; co < 8; ++co) {
              const size_t o = (((size_t)n * 8 + co) * a.H + oy) * a.W + ox;
              a.y[o] = yv[co]; a.cout[o] = cv[co];
            }
          }
          if (a.out_n8 || a.out_up) {
            const N8Px p = n8_make(xc, cv);
            if (a.out_n8) n8_put(a.out_n8, gs, n, gs.pixel(oy, ox), p);
            if (a.out_up) {
#pragma unroll
              for (int i = 0; i < 2; ++i)
#pragma unroll
                for (int j = 0; j < 2; ++j) n8_put(a.out_up, gu, n, gu.pixel(2 * oy + i, 2 * ox + j), p);
            }
          }
        }
        if (a.out_pool && !(a.dbg & 2)) {
          // The two independent 2x2/2 max-pools (step1.py:62-74): max over the window, NaN if any element is NaN — the
          // value ATen's `(v > m) || isnan(v)` chain leaves.  The two rows of a window belong to the warps of quarters
          // 2j and 2j+1: the even warp pools channels 0-3 and the odd warp channels 4-7, each handing the other its half
          // of the channels through shared memory (8 floats per lane, one 64-thread barrier per M-tile), and each writes
          // its 8-byte half of the pooled pixel's four N8 vectors.
          const int half = q & 1;                                   // channels [4*half, 4*half + 4) are pooled here
          float(*px)[8][32] = reinterpret_cast<float(*)[8][32]>(s_px + (size_t)(set * 2 + ((it / (uint32_t)a.esets) & 1)) * 4 * 8 * 32);
#pragma unroll
          for (int j = 0; j < 4; ++j) { px[q][j][lane] = half ? yv[j] : yv[4 + j]; px[q][4 + j][lane] = half ? cv[j] : cv[4 + j]; }
          asm volatile("bar.sync %0, 64;" ::"r"(1 + set * 2 + (q >> 1)) : "memory");
          float yp[4], cp[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float ya = half ? yv[4 + j] : yv[j], ca = half ? cv[4 + j] : cv[j];   // this warp's row
            const float yb = px[q ^ 1][j][lane], cb = px[q ^ 1][4 + j][lane];  // the partner's row
            const float ym = max_nan(ya, yb), cm = max_nan(ca, cb);            // vertical, then horizontal with lane + 1
            yp[j] = max_nan(ym, __shfl_down_sync(0xffffffffu, ym, 1));
            cp[j] = max_nan(cm, __shfl_down_sync(0xffffffffu, cm, 1));
          }
          const int py = oy >> 1, pxo = ox >> 1;
          if (valid && (lane & 1) == 0 && py < gp.H && pxo < gp.W) {
            float xcp[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) xcp[j] = yp[j] * cp[j];
            uint2 xh, xl, ch, cl;
            split4_fast(xcp, xh, xl);
            split4_fast(cp, ch, cl);
            __nv_bfloat16* b0 = a.out_pool + gp.plane(n, 0, 0) + gp.pixel(py, pxo) + 4 * half;
            const size_t pe = gp.plane_elems();
            *reinterpret_cast<uint2*>(b0) = xh;
            *reinterpret_cast<uint2*>(b0 + pe) = xl;
            *reinterpret_cast<uint2*>(b0 + 2 * pe) = ch;
            *reinterpret_cast<uint2*>(b0 + 3 * pe) = cl;
          }
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512));
  }
}

// ---- weight packing: blocks [NROW][8] bf16, row = dx*8 + co, column = input channel within the source --------------
//   block (src*K + dy)*2 + part (hi | lo), then one zero block
__global__ void nconv_tc_pack_kernel(const float* __restrict__ w, __nv_bfloat16* __restrict__ wp, int K, int NSRC, int NROW, int a_first) {
  const int nblk = 2 * NSRC * K + 1;
  const int total = nblk * NROW * 8;
  const int CI = 8 * NSRC;
  for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += gridDim.x * blockDim.x) {
    const int e = idx & 7, nrow = (idx >> 3) % NROW, blk = idx / (8 * NROW);
    float v = 0.f;
    const int part = blk & 1, sd = blk >> 1;
    const int dx = nrow >> 3, co = nrow & 7;
    if (blk < nblk - 1 && dx < K) {
      const int src = sd / K, dy = sd % K;
      // source A holds the first 8 weight input channels when a_first (or alone), else the last 8 (step1.py:80,85 vs :90)
      const int ciw = (NSRC == 1) ? e : ((src == 0) == (a_first != 0) ? e : 8 + e);
      v = __ldg(w + (((size_t)co * CI + ciw) * K + dy) * K + dx);
    }
    __nv_bfloat16 hi, lo;
    split_bf16(v, hi, lo);
    wp[idx] = part == 0 ? hi : lo;
  }
}

// ---- one launch for all ring fills / all weight packs of a DNET forward (dnet_pipeline.cu) -------------------------
__global__ void n8_zero_rings_multi_kernel(const N8RingJobs jobs) {
  const int j = blockIdx.y;
  if (j >= jobs.n) return;
  const N8Geom g = N8Geom::make(jobs.B, jobs.H[j], jobs.W[j]);
  __nv_bfloat16* buf = reinterpret_cast<__nv_bfloat16*>(jobs.buf[j]);
  const int ring_px = 2 * N8_RING * g.Wp + 2 * N8_RING * g.H;
  const long long total = (long long)g.B * 4 * ring_px;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
    const int r = (int)(idx % ring_px);
    const long long pl = idx / ring_px;
    int y, x;
    if (r < 2 * N8_RING * g.Wp) {
      const int rr = r / g.Wp; x = r % g.Wp;
      y = rr < N8_RING ? rr : g.Hp - 2 * N8_RING + rr;
    } else {
      const int q = r - 2 * N8_RING * g.Wp;
      const int cc = q % (2 * N8_RING);
      y = N8_RING + q / (2 * N8_RING);
      x = cc < N8_RING ? cc : g.Wp - 2 * N8_RING + cc;
    }
    *reinterpret_cast<uint4*>(buf + pl * g.plane_elems() + ((size_t)y * g.Wp + x) * 8) = make_uint4(0, 0, 0, 0);
  }
}

int n8_zero_rings_multi(const N8RingJobs& jobs, cudaStream_t st) {
  LaunchScope ls(st, "n8_zero_rings", 0.0, 0.0);
  n8_zero_rings_multi_kernel<<<dim3(148, jobs.n), 256, 0, st>>>(jobs);
  return (int)cudaGetLastError();
}

__global__ void nconv_tc_pack_multi_kernel(const NtPackJobs jobs) {
  const int j = blockIdx.y;
  if (j >= jobs.n) return;
  const int K = jobs.K[j], NSRC = jobs.nsrc[j], a_first = jobs.a_first[j];
  const int NROW = (K * 8 + 15) / 16 * 16;
  const float* w = jobs.w[j];
  __nv_bfloat16* wp = reinterpret_cast<__nv_bfloat16*>(jobs.wpack[j]);
  const int nblk = 2 * NSRC * K + 1, total = nblk * NROW * 8, CI = 8 * NSRC;
  for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += gridDim.x * blockDim.x) {
    const int e = idx & 7, nrow = (idx >> 3) % NROW, blk = idx / (8 * NROW);
    float v = 0.f;
    const int part = blk & 1, sd = blk >> 1;
    const int dx = nrow >> 3, co = nrow & 7;
    if (blk < nblk - 1 && dx < K) {
      const int src = sd / K, dy = sd % K;
      const int ciw = (NSRC == 1) ? e : ((src == 0) == (a_first != 0) ? e : 8 + e);
      v = __ldg(w + (((size_t)co * CI + ciw) * K + dy) * K + dx);
    }
    __nv_bfloat16 hi, lo;
    split_bf16(v, hi, lo);
    wp[idx] = part == 0 ? hi : lo;
  }
}

int nconv_tc_pack_multi(const NtPackJobs& jobs, cudaStream_t st) {
  LaunchScope ls(st, "tc_nconv_pack_weights", 0.0, 0.0);
  nconv_tc_pack_multi_kernel<<<dim3(4, jobs.n), 256, 0, st>>>(jobs);
  return (int)cudaGetLastError();
}

template <int K, int NSRC>
static int launch_nconv_tc(const nconv_b200_nconv_tc_desc* d, cudaStream_t st) {
  using C = NtCfg<K>;
  auto kern = nconv_tc_kernel<K, NSRC>;
  constexpr int NBLK = 2 * NSRC * K + 1;
  constexpr size_t W_BYTES = (NBLK * C::WBLK + 127) / 128 * 128;
  static bool attr_done[NCONV_MAX_DEVICES] = {};
  static int sm_count[NCONV_MAX_DEVICES] = {};
  int dev = 0, sms = 0;
  NCONV_RETURN_IF_ERR(cudaGetDevice(&dev));
  if (dev >= NCONV_MAX_DEVICES || !attr_done[dev]) {
    NCONV_RETURN_IF_ERR(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, NT_SMEM_MAX));
    if (dev < NCONV_MAX_DEVICES) attr_done[dev] = true;
  }
  if (dev < NCONV_MAX_DEVICES) sms = sm_count[dev];
  if (sms == 0) {
    NCONV_RETURN_IF_ERR(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    if (dev < NCONV_MAX_DEVICES) sm_count[dev] = sms;
  }
  // M-tiles (4 rows each) per tile: as many as two stages of NSRC*4 planes fit, capped by the image height
  int TM = (d->H + 3) / 4;
  const int cap = (K == 5) ? 10 : 6;
  if (TM > cap) TM = cap;
  while (TM > 1 && W_BYTES + NT_PX_BYTES + 2 * (size_t)NSRC * 4 * (4 * TM + C::HALO) * NT_ROW_BYTES > (size_t)NT_SMEM_MAX) --TM;
  NtArgs a = {};
  a.wpack = (const __nv_bfloat16*)d->wpack; a.w = d->w; a.bias = d->bias;
  a.out_n8 = (__nv_bfloat16*)d->out_n8; a.out_pool = (__nv_bfloat16*)d->out_pool; a.out_up = (__nv_bfloat16*)d->out_up;
  a.y = d->y; a.cout = d->cout; a.w_tail = d->w_tail; a.b_tail = d->b_tail; a.out_tail = d->out_tail;
  a.B = d->B; a.H = d->H; a.W = d->W; a.CI = 8 * NSRC;
  a.TM = TM; a.rows_ld = 4 * TM + C::HALO;
  a.tiles_x = (d->W + C::WT - 1) / C::WT; a.tiles_y = (d->H + 4 * TM - 1) / (4 * TM);
  a.ntiles = a.tiles_x * a.tiles_y * d->B;
  a.eps = d->eps;
  {
    static const char* ei = getenv("NCONV_B200_NT_ISSUERS");     // A/B measurements and debugging only
    static const char* es = getenv("NCONV_B200_NT_ESETS");
    a.issuers = ei && atoi(ei) == 1 ? 1 : NT_ISSUERS;
    a.esets = es && atoi(es) == 1 ? 1 : NT_ESETS;
    static const char* db = getenv("NCONV_B200_NT_DEBUG");
    a.dbg = db ? atoi(db) : 0;
  }
  size_t smem = W_BYTES + NT_PX_BYTES + 2 * (size_t)NSRC * 4 * a.rows_ld * NT_ROW_BYTES;
  if (smem < (size_t)NT_SMEM_MIN) smem = NT_SMEM_MIN;
  // N8 source seen as [planes = 4B][rows = H+4][(W+4)*4 32-bit words]; box [128 words][rows_ld][1]
  alignas(64) CUtensorMap tm[2];
  const void* srcs[2] = {d->a, d->b ? d->b : d->a};
  for (int si = 0; si < 2; ++si) {
    const int rc = tc_make_tmap(&tm[si], srcs[si], d->B * 4, d->H + 2 * N8_RING, d->W + 2 * N8_RING, NT_P, a.rows_ld);
    if (rc) return rc;
  }
  const int grid = a.ntiles < sms ? a.ntiles : sms;
  const double px = (double)d->B * d->H * d->W;
  double bytes = 64.0 * NSRC * px;
  if (d->out_n8) bytes += 64.0 * px;
  if (d->out_pool) bytes += 16.0 * px;
  if (d->out_up) bytes += 256.0 * px;
  if (d->y) bytes += 64.0 * px;
  if (d->out_tail) bytes += 4.0 * px;
  char label[64];
  snprintf(label, sizeof(label), "tc_nconv 8+%d->8 k%d @%dx%d", NSRC == 2 ? 8 : 0, K, d->H, d->W);
  {
    LaunchScope ls(st, label, 4.0 * 8 * NSRC * 8 * K * K * px, bytes);
    kern<<<grid, NT_THREADS, smem, st>>>(a, tm[0], tm[1]);
  }
  return (int)cudaGetLastError();
}

}  // namespace nconv

using namespace nconv;

extern "C" int nconv_b200_debug_buffer(void* pinned_host_64_bytes) {
  unsigned int* p = (unsigned int*)pinned_host_64_bytes;
  return (int)cudaMemcpyToSymbol(g_nt_dbg, &p, sizeof(p));
}

extern "C" size_t nconv_b200_n8_bytes(int B, int H, int W) {
  if (B <= 0 || H <= 0 || W <= 0) return 0;
  return N8Geom::make(B, H, W).total_elems() * sizeof(__nv_bfloat16) + N8_TAIL_SLACK_BYTES;
}

extern "C" int nconv_b200_n8_zero_ring(void* n8, int B, int H, int W, void* stream) {
  if (!n8 || B <= 0 || H <= 0 || W <= 0) return NCONV_ERR_BAD_ARG;
  const N8Geom g = N8Geom::make(B, H, W);
  const long long total = (long long)B * 4 * (2 * N8_RING * g.Wp + 2 * N8_RING * g.H);
  int blocks = (int)((total + 255) / 256);
  if (blocks > 148 * 8) blocks = 148 * 8;
  {
    LaunchScope ls((cudaStream_t)stream, "n8_zero_ring", 0.0, 16.0 * total);
    n8_zero_ring_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>((__nv_bfloat16*)n8, g);
  }
  return (int)cudaGetLastError();
}

extern "C" int nconv_b200_n8_pack(const float* x, const float* c, void* n8, int B, int H, int W, void* stream) {
  if (!x || !c || !n8 || B <= 0 || H <= 0 || W <= 0) return NCONV_ERR_BAD_ARG;
  const N8Geom g = N8Geom::make(B, H, W);
  const long long total = (long long)B * H * W;
  int blocks = (int)((total + 255) / 256);
  if (blocks > 148 * 16) blocks = 148 * 16;
  {
    LaunchScope ls((cudaStream_t)stream, "n8_pack", 0.0, 128.0 * total);
    n8_pack_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(x, c, (__nv_bfloat16*)n8, g);
  }
  return (int)cudaGetLastError();
}

extern "C" int nconv_b200_n8_unpack(const void* n8, float* xc, float* c, int B, int H, int W, void* stream) {
  if (!n8 || (!xc && !c) || B <= 0 || H <= 0 || W <= 0) return NCONV_ERR_BAD_ARG;
  const N8Geom g = N8Geom::make(B, H, W);
  const long long total = (long long)B * H * W;
  int blocks = (int)((total + 255) / 256);
  if (blocks > 148 * 16) blocks = 148 * 16;
  {
    LaunchScope ls((cudaStream_t)stream, "n8_unpack", 0.0, 128.0 * total);
    n8_unpack_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)n8, xc, c, g);
  }
  return (int)cudaGetLastError();
}

extern "C" int nconv_b200_nconv1_n8(const float* S, const float* w, const float* bias, void* out_n8, int B, int H, int W,
                                    float thr, float eps, void* stream) {
  if (!S || !w || !bias || !out_n8 || B <= 0 || H <= 0 || W <= 0) return NCONV_ERR_BAD_ARG;
  const N8Geom g = N8Geom::make(B, H, W);
  dim3 grid((W + N1_TW - 1) / N1_TW, (H + N1_TH - 1) / N1_TH, B);
  const double px = (double)B * H * W;
  {
    LaunchScope ls((cudaStream_t)stream, "nconv1_n8 1->8 k5", 4.0 * 8 * 25 * px, 68.0 * px);
    nconv1_n8_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(S, w, bias, (__nv_bfloat16*)out_n8, g, thr, eps);
  }
  return (int)cudaGetLastError();
}

extern "C" size_t nconv_b200_nconv_tc_wpack_bytes(int K, int CB) {
  if (!((K == 5 && CB == 0) || (K == 3 && CB == 8))) return 0;
  const int nsrc = CB ? 2 : 1, nrow = (K * 8 + 15) / 16 * 16;
  return (size_t)(2 * nsrc * K + 1) * nrow * 16;
}

extern "C" int nconv_b200_nconv_tc_pack_weights(const float* w, void* wpack, int K, int CB, int a_first, void* stream) {
  if (!w || !wpack) return NCONV_ERR_BAD_ARG;
  if (!((K == 5 && CB == 0) || (K == 3 && CB == 8))) return NCONV_ERR_UNSUPPORTED;
  const int nsrc = CB ? 2 : 1, nrow = (K * 8 + 15) / 16 * 16;
  const int total = (2 * nsrc * K + 1) * nrow * 8;
  {
    LaunchScope ls((cudaStream_t)stream, "tc_nconv_pack_weights", 0.0, 6.0 * total);
    nconv_tc_pack_kernel<<<(total + 255) / 256, 256, 0, (cudaStream_t)stream>>>(w, (__nv_bfloat16*)wpack, K, nsrc, nrow, a_first);
  }
  return (int)cudaGetLastError();
}

extern "C" int nconv_b200_nconv_tc(const nconv_b200_nconv_tc_desc* d, void* stream) {
  if (!d || !d->a || !d->wpack || !d->w || !d->bias) return NCONV_ERR_BAD_ARG;
  if (d->B <= 0 || d->H <= 0 || d->W <= 0) return NCONV_ERR_BAD_ARG;
  if ((d->y == nullptr) != (d->cout == nullptr)) return NCONV_ERR_BAD_ARG;
  if (d->out_tail && (!d->w_tail || !d->b_tail)) return NCONV_ERR_BAD_ARG;
  if (!d->out_n8 && !d->out_pool && !d->out_up && !d->y && !d->out_tail) return NCONV_ERR_BAD_ARG;
  if (d->out_pool && (d->H < 2 || d->W < 2)) return NCONV_ERR_BAD_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  if (d->K == 5 && d->CB == 0 && !d->b) return launch_nconv_tc<5, 1>(d, st);
  if (d->K == 3 && d->CB == 8 && d->b) return launch_nconv_tc<3, 2>(d, st);
  return NCONV_ERR_UNSUPPORTED;
}
