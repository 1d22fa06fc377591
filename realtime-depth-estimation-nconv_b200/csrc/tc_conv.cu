// tcgen05 implicit-GEMM 3x3 / stride-1 / pad-1 convolution for the dense layers of the RGB-guided
// stack (/root/reference/models/step2.py: ConvBlock :290-297, Basic2d :178-195 — the 19 layers
// with Ci in {32,64,128} and Co in {32,64} that hold 92 % of SETP2's flops).
//
//   y = act(scale[o] * conv3x3(cat[a, b]) + shift[o])            (scale/shift: folded eval BN or bias)
//
// Arithmetic: bf16x3 split precision.  Activations live in P8 (p8.cuh: hi/lo bf16, 16 mantissa
// bits), weights are split the same way, and every K=16 slice issues three tcgen05.mma
// (hi*hi + lo*hi + hi*lo) into one fp32 TMEM accumulator: ~2^-17 relative per product, i.e. fp32-class
// results (parity bound 1e-4, measured in tests/test_gpu_parity.py) at 1/3 of the bf16 tensor rate.
//
// Structure (one persistent CTA per SM, 192 threads, warp-specialised):
//   warp 0     producer: per K-stage (16 input channels) bulk-copies (cp.async.bulk, TMA engine,
//              mbarrier complete_tx) the halo tile rows of the 2x2 (hi|lo x chunk) planes and the
//              stage's packed weights into a 2-deep shared-memory ring
//   warp 1     one thread issues tcgen05.mma.cta_group::1.kind::f16 (M=128 pixels, N=Co, K=16):
//              the CTA tile is R rows x Wt columns of outputs, laid out FLAT in shared memory with
//              row pitch P = Wt+2, so 128 consecutive staged pixels are one A tile and tap (u,v) is
//              the descriptor start address + (u*P+v)*16 B.  Halo columns produce garbage rows
//              that the epilogue drops (2/P of the MMA rows).
//   warps 2-5  epilogue: tcgen05.ld the fp32 accumulators (double-buffered in TMEM so the next
//              tile's MMAs overlap), scale/shift/ReLU, split to hi/lo bf16, 16-byte coalesced
//              stores into the P8 output.
#include "common.cuh"
#include "launch_count.h"
#include "nconv_b200.h"
#include "p8.cuh"
#include <cstdio>

namespace nconv {

constexpr int TC_THREADS = 192;
constexpr int TC_STAGES = 2;
constexpr int TC_SMEM_MAX = 227 * 1024 - 1024;   // dynamic part; static barriers take the rest

struct TcArgs {
  const __nv_bfloat16* src[2];  // P8 sources (src[1] may be null)
  int c8[2];                    // chunks (planes/2) per image of each source
  int pairs[2];                 // 16-channel K-stages taken from each source
  const __nv_bfloat16* wpack;   // [kstage][tap][part][j][CO][8]
  const float* scale;           // [CO] or null
  const float* shift;           // [CO] or null
  __nv_bfloat16* out;           // P8 [B][CO/8][2][Hp][Wp][8]
  int B, H, W, Hp, Wp;
  int R, Wt, P, T, rows_alloc;  // tile: R x Wt outputs, pitch P, T M-tiles of 128, staged rows
  int tiles_x, tiles_y, ntiles;
  int relu;
};

// ---- PTX wrappers -----------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ uint32_t mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.b32 %0, 1, 0, p;\n\t}"
      : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
  return ok;
}
// Bounded wait: a protocol bug must fault (trap -> launch error), never hang the GPU.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t spins = 0;
  while (!mbar_try_wait(bar, parity)) {
    if (++spins > (1u << 22)) __trap();
  }
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void umma_bf16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                          uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// K-major, no-swizzle shared-memory matrix descriptor (cute::UMMA::SmemDescriptor bit layout):
// rows of 16 B, 8-row core matrices `sbo` bytes apart, the two K-chunks of a K=16 slice `lbo` apart.
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  return (uint64_t)((saddr >> 4) & 0x3FFFu) | ((uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16) |
         ((uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32) | ((uint64_t)1 << 46);
}

template <int N>
__global__ void __launch_bounds__(TC_THREADS, 1) conv3x3_tc_kernel(const TcArgs a) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t s_full[TC_STAGES], s_empty[TC_STAGES], s_tfull[2], s_tempty[2];
  __shared__ uint32_t s_tmem_base;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t a_plane_bytes = (uint32_t)a.rows_alloc * a.P * 16;       // one (part, j) plane
  const uint32_t a_stage_bytes = 4 * a_plane_bytes;
  constexpr uint32_t B_TAP_BYTES = 2 * 2 * N * 16;                        // [part][j][N][16 B]
  constexpr uint32_t B_STAGE_BYTES = 9 * B_TAP_BYTES;
  const uint32_t stage_bytes = a_stage_bytes + B_STAGE_BYTES;
  const uint32_t smem_base = smem_u32(smem);
  const int nk = a.pairs[0] + a.pairs[1];

  if (threadIdx.x == 0) {
    for (int s = 0; s < TC_STAGES; ++s) { mbar_init(smem_u32(&s_full[s]), 1); mbar_init(smem_u32(&s_empty[s]), 1); }
    for (int b = 0; b < 2; ++b) { mbar_init(smem_u32(&s_tfull[b]), 1); mbar_init(smem_u32(&s_tempty[b]), 4); }
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

  if (warp == 0) {
    // ===================== producer =====================
    uint32_t stage = 0, phase = 0;
    for (int tile = blockIdx.x; tile < a.ntiles; tile += gridDim.x) {
      const int tx = tile % a.tiles_x, ty = (tile / a.tiles_x) % a.tiles_y, n = tile / (a.tiles_x * a.tiles_y);
      const int y0 = ty * a.R, x0 = tx * a.Wt;                 // padded coords of the first staged row / col
      int nrow = a.R + 2;
      if (y0 + nrow > a.Hp) nrow = a.Hp - y0;
      const uint32_t row_bytes = (uint32_t)a.P * 16;
      for (int k = 0; k < nk; ++k) {
        const int si = (k < a.pairs[0]) ? 0 : 1;
        const int c0 = 2 * (si == 0 ? k : k - a.pairs[0]);
        mbar_wait(smem_u32(&s_empty[stage]), phase ^ 1);
        const uint32_t full = smem_u32(&s_full[stage]);
        const uint32_t sa = smem_base + stage * stage_bytes;
        if (lane == 0) {
          mbar_arrive_expect_tx(full, 4u * nrow * row_bytes + B_STAGE_BYTES);
          bulk_g2s(sa + a_stage_bytes, a.wpack + (size_t)k * (B_STAGE_BYTES / 2), B_STAGE_BYTES, full);
        }
        __syncwarp();
        const size_t plane_elems = (size_t)a.Hp * a.Wp * 8;
        for (int i = lane; i < 4 * nrow; i += 32) {
          const int pj = i / nrow, r = i - pj * nrow;          // pj = part*2 + j
          const int part = pj >> 1, j = pj & 1;
          const __nv_bfloat16* g = a.src[si] + (((size_t)n * a.c8[si] + c0 + j) * 2 + part) * plane_elems +
                                   ((size_t)(y0 + r) * a.Wp + x0) * 8;
          bulk_g2s(sa + pj * a_plane_bytes + r * row_bytes, g, row_bytes, full);
        }
        __syncwarp();
        if (++stage == TC_STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    constexpr uint32_t IDESC = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((128u >> 4) << 24);
    uint32_t stage = 0, phase = 0, it = 0;
    for (int tile = blockIdx.x; tile < a.ntiles; tile += gridDim.x, ++it) {
      const uint32_t buf = it & 1, tphase = (it >> 1) & 1;
      mbar_wait(smem_u32(&s_tempty[buf]), tphase ^ 1);
      tc_fence_after();
      for (int k = 0; k < nk; ++k) {
        mbar_wait(smem_u32(&s_full[stage]), phase);
        tc_fence_after();
        if (lane == 0) {
          const uint32_t sa = smem_base + stage * stage_bytes;
          const uint32_t sb = sa + a_stage_bytes;
          for (int t = 0; t < a.T; ++t) {
            const uint32_t d = tmem_base + buf * 256 + t * N;
            uint32_t acc = (k > 0) ? 1u : 0u;
#pragma unroll
            for (int tap = 0; tap < 9; ++tap) {
              const int u = tap / 3, v = tap - 3 * u;
              const uint32_t aoff = (uint32_t)(u * a.P + v + t * 128) * 16;
              const uint64_t a_hi = umma_desc(sa + aoff, a_plane_bytes, 128);
              const uint64_t a_lo = umma_desc(sa + 2 * a_plane_bytes + aoff, a_plane_bytes, 128);
              const uint64_t b_hi = umma_desc(sb + tap * B_TAP_BYTES, N * 16, 128);
              const uint64_t b_lo = umma_desc(sb + tap * B_TAP_BYTES + 2 * N * 16, N * 16, 128);
              umma_bf16(d, a_hi, b_hi, IDESC, acc);
              umma_bf16(d, a_lo, b_hi, IDESC, 1u);
              umma_bf16(d, a_hi, b_lo, IDESC, 1u);
              acc = 1u;
            }
          }
          umma_commit(smem_u32(&s_empty[stage]));                       // frees the smem slot when the MMAs retire
          if (k == nk - 1) umma_commit(smem_u32(&s_tfull[buf]));       // accumulators of this tile complete
        }
        __syncwarp();
        if (++stage == TC_STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else {
    // ===================== epilogue (warps 2..5) =====================
    const int q = warp & 3;                                    // TMEM lane quarter this warp may access
    uint32_t it = 0;
    const size_t plane_elems = (size_t)a.Hp * a.Wp * 8;
    for (int tile = blockIdx.x; tile < a.ntiles; tile += gridDim.x, ++it) {
      const uint32_t buf = it & 1, tphase = (it >> 1) & 1;
      const int tx = tile % a.tiles_x, ty = (tile / a.tiles_x) % a.tiles_y, n = tile / (a.tiles_x * a.tiles_y);
      const int y0 = ty * a.R, x0 = tx * a.Wt;
      mbar_wait(smem_u32(&s_tfull[buf]), tphase);
      tc_fence_after();
      for (int t = 0; t < a.T; ++t) {
        const int m = t * 128 + q * 32 + lane;                 // flat staged position of this thread's pixel
        const int r = m / a.P, c = m - r * a.P;
        const int oy = y0 + r, ox = x0 + c;
        const bool valid = (r < a.R) && (c < a.Wt) && (oy < a.H) && (ox < a.W);
        __nv_bfloat16* obase = a.out + ((size_t)n * (N / 8) * 2) * plane_elems + ((size_t)(oy + 1) * a.Wp + (ox + 1)) * 8;
#pragma unroll
        for (int cb = 0; cb < N / 32; ++cb) {
          uint32_t v[32];
          tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + buf * 256 + t * N + cb * 32, v);
          tmem_ld_wait();
          if (valid) {
#pragma unroll
            for (int ch = 0; ch < 4; ++ch) {
              float f[8];
#pragma unroll
              for (int e = 0; e < 8; ++e) {
                const int o = cb * 32 + ch * 8 + e;
                float x = __uint_as_float(v[ch * 8 + e]);
                const float sc = a.scale ? __ldg(a.scale + o) : 1.f;
                const float sh = a.shift ? __ldg(a.shift + o) : 0.f;
                x = fmaf(x, sc, sh);
                f[e] = a.relu ? fmaxf(x, 0.f) : x;
              }
              uint4 hi, lo;
              split8(f, hi, lo);
              __nv_bfloat16* p = obase + (size_t)(cb * 4 + ch) * 2 * plane_elems;
              *reinterpret_cast<uint4*>(p) = hi;
              *reinterpret_cast<uint4*>(p + plane_elems) = lo;
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(smem_u32(&s_tempty[buf]));
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512));
  }
}

// ---- weight packing: w [CO][CI][3][3] fp32 -> [kstage][tap][part][j][CO][8] split bf16 -----------
__global__ void tc_pack_weights_kernel(const float* __restrict__ w, __nv_bfloat16* __restrict__ wp, int CO, int CI) {
  const int total = (CI / 16) * 9 * 2 * CO * 8;   // (kstage, tap, j, co, e); both parts written together
  for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += gridDim.x * blockDim.x) {
    int e = idx & 7;
    int r = idx >> 3;
    int co = r % CO; r /= CO;
    int j = r & 1; r >>= 1;
    int tap = r % 9;
    int ks = r / 9;
    const int ci = ks * 16 + j * 8 + e;
    __nv_bfloat16 hi, lo;
    split_bf16(__ldg(w + ((size_t)co * CI + ci) * 9 + tap), hi, lo);
    const size_t base = ((size_t)(ks * 9 + tap) * 2) * 2 * CO * 8;     // [part][j][CO][8]
    wp[base + ((size_t)(0 * 2 + j) * CO + co) * 8 + e] = hi;
    wp[base + ((size_t)(1 * 2 + j) * CO + co) * 8 + e] = lo;
  }
}

// ---- P8 <-> NCHW fp32 (boundaries of the stack and tests) ----------------------------------------
__global__ void p8_pack_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ p8, int B, int C, int H, int W) {
  const P8Geom g = P8Geom::make(B, C, H, W);
  const size_t total = (size_t)B * g.C8 * g.Hp * g.Wp;
  for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
    int xx = (int)(idx % g.Wp);
    size_t r = idx / g.Wp;
    int yy = (int)(r % g.Hp); r /= g.Hp;
    int ch = (int)(r % g.C8);
    int n = (int)(r / g.C8);
    float f[8];
    const bool inside = yy >= 1 && yy <= H && xx >= 1 && xx <= W;
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int c = ch * 8 + e;
      f[e] = (inside && c < C) ? __ldg(x + (((size_t)n * C + c) * H + (yy - 1)) * W + (xx - 1)) : 0.f;
    }
    uint4 hi, lo;
    split8(f, hi, lo);
    const size_t off = ((size_t)yy * g.Wp + xx) * 8;
    *reinterpret_cast<uint4*>(p8 + g.plane(n, ch, 0) + off) = hi;
    *reinterpret_cast<uint4*>(p8 + g.plane(n, ch, 1) + off) = lo;
  }
}

__global__ void p8_unpack_kernel(const __nv_bfloat16* __restrict__ p8, float* __restrict__ x, int B, int C, int H, int W) {
  const P8Geom g = P8Geom::make(B, C, H, W);
  const size_t total = (size_t)B * g.C8 * H * W;
  for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
    int xx = (int)(idx % W);
    size_t r = idx / W;
    int yy = (int)(r % H); r /= H;
    int ch = (int)(r % g.C8);
    int n = (int)(r / g.C8);
    const size_t off = g.pixel(yy, xx);
    const uint4 hi = *reinterpret_cast<const uint4*>(p8 + g.plane(n, ch, 0) + off);
    const uint4 lo = *reinterpret_cast<const uint4*>(p8 + g.plane(n, ch, 1) + off);
    float f[8];
    join8(hi, lo, f);
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int c = ch * 8 + e;
      if (c < C) x[(((size_t)n * C + c) * H + yy) * W + xx] = f[e];
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

// ---- tile geometry ---------------------------------------------------------------------------------
struct TcPlan { int R, Wt, P, T, rows_alloc, tiles_x, tiles_y; size_t smem; bool ok; };

static TcPlan tc_plan(int H, int W, int N) {
  TcPlan best = {}; best.ok = false;
  double best_cost = 1e300;
  const int tmax = 256 / N;                            // accumulator columns per TMEM buffer
  for (int Wt = 16; Wt <= 254 && Wt <= W + 15; ++Wt) {
    const int P = Wt + 2;
    for (int R = 1; R <= 32; ++R) {
      const int flat = R * P - 2;
      const int T = (flat + 127) / 128;
      if (T > tmax) break;
      const int rows_alloc = (T * 128 + 2 * P + 2 + P - 1) / P;
      const size_t smem = (size_t)TC_STAGES * (4ull * rows_alloc * P * 16 + 9ull * 2 * 2 * N * 16) + 256;
      if (smem > (size_t)TC_SMEM_MAX) continue;
      const int tx = (W + Wt - 1) / Wt, ty = (H + R - 1) / R;
      // MMA rows issued, plus a load term (halo rows re-read), in arbitrary but comparable units
      const double cost = (double)tx * ty * (T * 128.0 + 0.35 * (R + 2) * P);
      if (cost < best_cost) {
        best_cost = cost;
        best = {R, Wt, P, T, rows_alloc, tx, ty, smem, true};
      }
    }
  }
  return best;
}

template <int N>
static int launch_tc(TcArgs a, const TcPlan& p, cudaStream_t st) {
  auto kern = conv3x3_tc_kernel<N>;
  static int max_smem_set[NCONV_MAX_DEVICES] = {};
  int dev = 0;
  NCONV_RETURN_IF_ERR(cudaGetDevice(&dev));
  if (dev >= NCONV_MAX_DEVICES || max_smem_set[dev] < (int)p.smem) {
    NCONV_RETURN_IF_ERR(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM_MAX));
    if (dev < NCONV_MAX_DEVICES) max_smem_set[dev] = TC_SMEM_MAX;
  }
  static int sm_count[NCONV_MAX_DEVICES] = {};
  int sms = (dev < NCONV_MAX_DEVICES) ? sm_count[dev] : 0;
  if (sms == 0) {
    NCONV_RETURN_IF_ERR(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    if (dev < NCONV_MAX_DEVICES) sm_count[dev] = sms;
  }
  const int grid = a.ntiles < sms ? a.ntiles : sms;
  const int CI = 16 * (a.pairs[0] + a.pairs[1]);
  const double px = (double)a.B * a.H * a.W;
  char label[64];
  snprintf(label, sizeof(label), "tc_conv3x3 %d->%d @%dx%d", CI, N, a.H, a.W);
  {
    // algorithmic: fp32-equivalent activations in and out (P8 is the same 4 B per value)
    LaunchScope ls(st, label, 2.0 * CI * N * 9 * px, 4.0 * (CI + N) * px);
    kern<<<grid, TC_THREADS, p.smem, st>>>(a);
  }
  return (int)cudaGetLastError();
}

}  // namespace nconv

using namespace nconv;

extern "C" size_t nconv_b200_p8_bytes(int B, int C, int H, int W) {
  if (B <= 0 || C <= 0 || H <= 0 || W <= 0) return 0;
  return P8Geom::make(B, C, H, W).total_elems() * sizeof(__nv_bfloat16) + P8_TAIL_SLACK_BYTES;
}

extern "C" int nconv_b200_p8_pack(const float* x, void* p8, int B, int C, int H, int W, void* stream) {
  if (!x || !p8 || B <= 0 || C <= 0 || H <= 0 || W <= 0) return NCONV_ERR_BAD_ARG;
  const P8Geom g = P8Geom::make(B, C, H, W);
  const size_t total = (size_t)B * g.C8 * g.Hp * g.Wp;
  int blocks = (int)((total + 255) / 256); if (blocks > 148 * 16) blocks = 148 * 16;
  {
    LaunchScope ls((cudaStream_t)stream, "p8_pack", 0.0, 8.0 * B * C * H * W);
    p8_pack_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(x, (__nv_bfloat16*)p8, B, C, H, W);
  }
  return (int)cudaGetLastError();
}

extern "C" int nconv_b200_p8_unpack(const void* p8, float* x, int B, int C, int H, int W, void* stream) {
  if (!x || !p8 || B <= 0 || C <= 0 || H <= 0 || W <= 0) return NCONV_ERR_BAD_ARG;
  const P8Geom g = P8Geom::make(B, C, H, W);
  const size_t total = (size_t)B * g.C8 * H * W;
  int blocks = (int)((total + 255) / 256); if (blocks > 148 * 16) blocks = 148 * 16;
  {
    LaunchScope ls((cudaStream_t)stream, "p8_unpack", 0.0, 8.0 * B * C * H * W);
    p8_unpack_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)p8, x, B, C, H, W);
  }
  return (int)cudaGetLastError();
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

extern "C" size_t nconv_b200_tc_wpack_bytes(int CO, int CI) {
  if (CO <= 0 || CI <= 0 || CI % 16) return 0;
  return (size_t)(CI / 16) * 9 * 2 * 2 * CO * 8 * sizeof(__nv_bfloat16);
}

extern "C" int nconv_b200_tc_pack_weights(const float* w, void* wpack, int CO, int CI, void* stream) {
  if (!w || !wpack || CO <= 0 || CI <= 0 || CI % 16) return NCONV_ERR_BAD_ARG;
  const int total = (CI / 16) * 9 * 2 * CO * 8;
  {
    LaunchScope ls((cudaStream_t)stream, "tc_pack_weights", 0.0, 8.0 * CO * CI * 9);
    tc_pack_weights_kernel<<<(total + 255) / 256, 256, 0, (cudaStream_t)stream>>>(w, (__nv_bfloat16*)wpack, CO, CI);
  }
  return (int)cudaGetLastError();
}

extern "C" int nconv_b200_conv3x3_tc(const nconv_b200_tc_desc* d, void* stream) {
  if (!d || !d->a || !d->wpack || !d->y) return NCONV_ERR_BAD_ARG;
  if (d->B <= 0 || d->H <= 0 || d->W <= 0 || d->CA <= 0 || d->CB < 0) return NCONV_ERR_BAD_ARG;
  if (d->CA % 16 || d->CB % 16 || (d->CB > 0 && !d->b)) return NCONV_ERR_UNSUPPORTED;
  if (!(d->CO == 32 || d->CO == 64)) return NCONV_ERR_UNSUPPORTED;
  const TcPlan p = tc_plan(d->H, d->W, d->CO);
  if (!p.ok) return NCONV_ERR_UNSUPPORTED;
  TcArgs a = {};
  a.src[0] = (const __nv_bfloat16*)d->a; a.src[1] = (const __nv_bfloat16*)d->b;
  a.c8[0] = d->CA / 8; a.c8[1] = d->CB / 8;
  a.pairs[0] = d->CA / 16; a.pairs[1] = d->CB / 16;
  a.wpack = (const __nv_bfloat16*)d->wpack; a.scale = d->scale; a.shift = d->shift;
  a.out = (__nv_bfloat16*)d->y;
  a.B = d->B; a.H = d->H; a.W = d->W; a.Hp = d->H + 2; a.Wp = d->W + 2;
  a.R = p.R; a.Wt = p.Wt; a.P = p.P; a.T = p.T; a.rows_alloc = p.rows_alloc;
  a.tiles_x = p.tiles_x; a.tiles_y = p.tiles_y; a.ntiles = p.tiles_x * p.tiles_y * d->B;
  a.relu = d->relu;
  cudaStream_t st = (cudaStream_t)stream;
  return d->CO == 64 ? launch_tc<64>(a, p, st) : launch_tc<32>(a, p, st);
}
