// tcgen05 implicit-GEMM convolutions for the dense layers of the RGB-guided stack
// (/root/reference/models/step2.py): 98.8 % of SETP2's flops are convolutions with Ci,Co >= 32
// (SURVEY App. A table 2).  One kernel, three tap tables:
//
//   S1  3x3 / stride 1 / pad 1         ConvBlock :290-297, Basic2d :178-195 (two-source concat)
//   S2  3x3 / stride 2 / pad 1 (+ the 1x1 / stride 2 residual branch)      RGBEncoder :134-154
//   T   ConvTranspose2d 4x4 / stride 2 / pad 1 as four sub-pixel 2x2 convolutions
//                                                                          Basic2dTrans :197-214
//
//   y = act(scale[o] * conv(cat[a, b]) + shift[o]) (+ conv1x1_s2(cat[a, b]) for S2)
//
// Arithmetic: bf16x3 split precision.  Activations live in P8 / P8Q (p8.cuh: hi/lo bf16, 16 mantissa
// bits), weights are split the same way, and every K=16 slice issues three tcgen05.mma
// (hi*hi + lo*hi + hi*lo) into one fp32 TMEM accumulator: ~2^-17 relative per product, i.e.
// fp32-class results (parity bound 1e-4, measured in tests/test_gpu_parity.py) at 1/3 of the bf16
// tensor rate.
//
// Structure (one persistent CTA per SM, 352 threads, warp-specialised):
//   warp 0     producer: per K-stage (16 input channels) bulk-copies (cp.async.bulk on the TMA
//              engine, mbarrier complete_tx) the halo rows of the staged planes and the stage's
//              packed weights into a 2-deep shared-memory ring
//   warp 1     one thread issues tcgen05.mma.cta_group::1.kind::f16 (M=128 anchors, N=Co, K=16).
//              The CTA tile is R rows x Wt columns of anchors (output pixels; input pixels for T)
//              staged FLAT with row pitch P = Wt+2, so 128 consecutive staged pixels are one
//              K-major no-swizzle A tile and a tap is "descriptor start address + shift*16 B"
//              (implicit im2col; for S2 the shift is taken inside one of the four phase planes).
//              Halo columns yield garbage rows that the epilogue drops (2/P of the MMA rows).
//   warps 2-5, 7-10  epilogue, one (Co = 64) or two (Co = 32) sets of four warps (one per TMEM lane quarter): tcgen05.ld the fp32
//              accumulators (double-buffered in TMEM so the next tile's MMAs overlap), scale/shift/ReLU
//              (+ residual accumulator), split to hi/lo bf16, 16-byte stores into the P8 output (and its
//              P8Q copy for a stride-2 consumer).  The (M-tile, phase, 32-channel block) work items of a
//              buffer alternate between the sets (TcEsets below).
#include "common.cuh"
#include "launch_count.h"
#include "nconv_b200.h"
#include "p8.cuh"
#include "tc_ptx.cuh"
#include "tc_tmap.cuh"
#include <cuda.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace nconv {

constexpr int TC_THREADS = 352;      // warp 0, 6: producers; warp 1: MMA issuer; warps 2-5 and (Co = 32) 7-10: the epilogue sets
// Epilogue sets per output width.  Measured A/B in one process (bench, B = 32): a second set makes the Co = 32 kernels
// 6 % faster (32->32 @480x640 0.748 -> 0.700 ms, transposed 48->32 -2 %); with Co = 64 the MMAs are the longer side and the
// extra warps only cost (+1 %, stride-2 +3 %), so those launch without them.
template <int N> struct TcEsets { static constexpr int value = (N == 32) ? 2 : 1; };
constexpr int TC_MAX_STAGES = 3;
constexpr int TC_SMEM_MAX = 227 * 1024 - 1024;   // dynamic part; static barriers etc. take the rest
constexpr int TC_MAX_TAPS = 17;

enum { TC_S1 = 0, TC_S2 = 1, TC_T = 2 };

// Compile-time tap tables (tap index -> staged-tile offset in 16 B units, weight block, accumulator
// group, first-into-group).  `plane16` = one staged plane in 16 B units; plane group g starts at
// g*4*plane16.
template <int MODE> struct TapTable;
template <> struct TapTable<TC_S1> {            // tap = u*3+v reads staged (r+u, c+v)
  static constexpr int NTAPS = 9;
  __host__ __device__ static constexpr uint32_t a_off16(int i, int P, uint32_t) { return (uint32_t)((i / 3) * P + (i % 3)); }
  __host__ __device__ static constexpr uint32_t w_idx(int i) { return (uint32_t)i; }
  __host__ __device__ static constexpr uint32_t dgrp(int) { return 0; }
  __host__ __device__ static constexpr bool first(int i) { return i == 0; }
};
template <> struct TapTable<TC_S2> {
  // input row 2y+u-1: u=0 -> odd phase, phase row y-1 (staged row r); u=1 -> even phase, row y
  // (staged r+1); u=2 -> odd phase, row y (staged r+1).  Same along x.  Tap 9 = the 1x1/s2 branch.
  static constexpr int NTAPS = 10;
  __host__ __device__ static constexpr uint32_t a_off16(int i, int P, uint32_t plane16) {
    if (i == 9) return (uint32_t)(P + 1);                               // x[2y, 2x]: phase (0,0)
    const int u = i / 3, v = i % 3;
    const int py = (u == 1) ? 0 : 1, rs = (u == 0) ? 0 : 1, px = (v == 1) ? 0 : 1, cs = (v == 0) ? 0 : 1;
    return (uint32_t)(py * 2 + px) * 4 * plane16 + (uint32_t)(rs * P + cs);
  }
  __host__ __device__ static constexpr uint32_t w_idx(int i) { return (uint32_t)i; }
  __host__ __device__ static constexpr uint32_t dgrp(int i) { return i == 9 ? 1 : 0; }
  __host__ __device__ static constexpr bool first(int i) { return i == 0 || i == 9; }
};
template <> struct TapTable<TC_T> {
  // out[2a+py, 2b+px] += in[a+py-t, b+px-s] * W[.., kh=(1-py)+2t, kw=(1-px)+2s];  tap = kh*4+kw
  static constexpr int NTAPS = 16;
  __host__ __device__ static constexpr int py(int i) { return 1 - ((i / 4) & 1); }
  __host__ __device__ static constexpr int px(int i) { return 1 - ((i % 4) & 1); }
  __host__ __device__ static constexpr uint32_t a_off16(int i, int P, uint32_t) {
    const int dy = py(i) - ((i / 4) >> 1), dx = px(i) - ((i % 4) >> 1);
    return (uint32_t)((1 + dy) * P + (1 + dx));
  }
  __host__ __device__ static constexpr uint32_t w_idx(int i) { return (uint32_t)i; }
  __host__ __device__ static constexpr uint32_t dgrp(int i) { return (uint32_t)(py(i) * 2 + px(i)); }
  // first tap of each phase in kh-major order: kh in {0,1} x kw in {0,1}
  __host__ __device__ static constexpr bool first(int i) { return (i / 4) < 2 && (i % 4) < 2; }
};

struct TcTap {
  uint32_t a_off;   // byte offset of the tap's hi A tile inside a stage (plane group base + shift*16)
  uint16_t w_idx;   // weight tap block inside the stage's packed weights
  uint8_t dgrp;     // accumulator group
  uint8_t first;    // first tap into its accumulator group (overwrites at K-stage 0)
};

struct TcArgs {
  const __nv_bfloat16* src[2];  // sources (P8, or P8Q for S2); src[1] may be null
  int c8[2];                    // chunks per image of each source (1: a lone channel, duplicated as both K-chunks)
  int pairs[2];                 // K-stages (16 channels) taken from each source
  size_t slab_elems;            // elements per (n, chunk, part) slab of a source
  size_t grp_elems;             // elements per plane group inside a slab (P8Q phase plane)
  int src_pitch, src_rows;      // source plane geometry (pixels per row, rows)
  const __nv_bfloat16* wpack;   // [kstage][wtap][j][part][CO][8]
  const float* scale;           // [CO] or null
  const float* shift;           // [CO] or null
  __nv_bfloat16* out;           // P8 [B][CO/8][2][OH+2][OW+2][8]
  __nv_bfloat16* outq;          // optional P8Q copy of the output
  float* out32;                 // S1 only: fp32 NCHW output instead of `out` (channels [o32_c0, o32_c0+CO) of a [B][o32_ctot][OH][OW] tensor)
  int o32_ctot, o32_c0;
  int o32_phase_c;              // > 0: out32 is [B][phase_c][2*OH][2*OW] and output channel (py*2+px)*phase_c + c lands at [c][2y+py][2x+px]
  int B, AH, AW, OH, OW;        // anchor grid and output size
  int R, Wt, P, T, rows_alloc, rows_ld, G, ngrp, ntaps, ntaps_w;
  uint32_t P_magic;             // ceil(2^32 / P): m / P == __umulhi(m, P_magic) for m * P < 2^32 (no integer division in the epilogue)
  int tiles_x, tiles_y, ntiles;
  int relu, mode, stages;
  int dbg;                      // NCONV_B200_TC_DEBUG bits (timing experiments only): 1 no stores, 2 no MMAs, 4 no loads, 64 one epilogue set
  int use_tmap;                 // halo planes by TMA tensor copies (one box per staged plane) instead of per-row bulk copies
  TcTap taps[TC_MAX_TAPS];
};

// STACK (Co = 32, S1): the packed weights of a tap are [j][hi|lo][N][8], so ONE B descriptor with 2N rows
// is [W_hi ; W_lo]: A_hi x [W_hi|W_lo] gives hi*hi and hi*lo in one N=2N MMA (two column halves of the
// accumulator, summed by the epilogue) and A_lo x W_hi is the second MMA.  An M128xK16 MMA reads its 4 KB A
// tile from shared memory in ~32 cycles whatever N is, so at N=32 the three MMAs of a slice were bound by
// three A reads (96 cycles); stacked they are two (64 cycles).
template <int N, int MODE, bool STACK>
__global__ void __launch_bounds__(TC_THREADS, 1) conv_tc_kernel(const __grid_constant__ TcArgs a, const __grid_constant__ CUtensorMap tm0,
                                                                const __grid_constant__ CUtensorMap tm1) {
  constexpr int NACC = STACK ? 2 * N : N;                                   // accumulator columns per (group, M-tile)
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t s_full[TC_MAX_STAGES], s_empty[TC_MAX_STAGES], s_tfull[2], s_tempty[2];
  __shared__ uint32_t s_tmem_base;
  __shared__ float s_scale[N], s_shift[N];

  // shuffle: provably warp-uniform warp index, so the MMA issuer's code compiles to the uniform datapath
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
  const uint32_t a_plane_bytes = (uint32_t)a.rows_alloc * a.P * 16;       // one (group, part, j) plane
  const uint32_t a_stage_bytes = (uint32_t)a.G * 4 * a_plane_bytes;
  constexpr uint32_t B_TAP_BYTES = 2 * 2 * N * 16;                        // [j][part][N][16 B]
  const uint32_t b_stage_bytes = (uint32_t)a.ntaps_w * B_TAP_BYTES;
  const uint32_t stage_bytes = a_stage_bytes + b_stage_bytes;
  const uint32_t smem_base = smem_u32(smem);
  const int nk = a.pairs[0] + a.pairs[1];

  if (threadIdx.x < N) {
    s_scale[threadIdx.x] = a.scale ? __ldg(a.scale + threadIdx.x) : 1.f;
    s_shift[threadIdx.x] = a.shift ? __ldg(a.shift + threadIdx.x) : 0.f;
  }
  if (threadIdx.x == 0) {
    for (int s = 0; s < TC_MAX_STAGES; ++s) { mbar_init(smem_u32(&s_full[s]), 1); mbar_init(smem_u32(&s_empty[s]), 1); }
    for (int b = 0; b < 2; ++b) { mbar_init(smem_u32(&s_tfull[b]), 1); mbar_init(smem_u32(&s_tempty[b]), 4 * TcEsets<N>::value); }
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

  if (warp == 0 || warp == 6) {
    // ===================== producers =====================
    // A bulk copy takes its operands from uniform registers, so a warp issues its lanes' copies one at a
    // time (~60 cycles each); with ~45 row copies per K-stage one warp could not keep ahead of the MMAs
    // (ncu: the issuer sat on the `full` barrier).  Two warps split the rows; warp 0 also posts the
    // expected byte count and the weight copy (the transaction count may go negative in between).
    const int pw = (warp == 0) ? 0 : 1;
    uint32_t stage = 0, phase = 0;
    const uint32_t row_bytes = (uint32_t)a.P * 16;
    for (int tile = blockIdx.x; tile < a.ntiles; tile += gridDim.x) {
      const int tx = tile % a.tiles_x, ty = (tile / a.tiles_x) % a.tiles_y, n = tile / (a.tiles_x * a.tiles_y);
      const int y0 = ty * a.R, x0 = tx * a.Wt;                 // plane coords of the first staged row / col
      int nrow = a.rows_ld;
      if (y0 + nrow > a.src_rows) nrow = a.src_rows - y0;
      const int ncopy = a.G * 4 * nrow;
      for (int k = 0; k < nk; ++k) {
        const int si = (k < a.pairs[0]) ? 0 : 1;
        const bool lone = a.c8[si] == 1;
        const int c0 = lone ? 0 : 2 * (si == 0 ? k : k - a.pairs[0]);
        mbar_wait(smem_u32(&s_empty[stage]), phase ^ 1);
        const uint32_t full = smem_u32(&s_full[stage]);
        const uint32_t sa = smem_base + stage * stage_bytes;
        if (a.dbg & 4) { if (pw == 0 && lane == 0) mbar_arrive(full); __syncwarp(); if (++stage == (uint32_t)a.stages) { stage = 0; phase ^= 1; } continue; }
        if (a.use_tmap) {
          // one tensor copy per staged plane: box [P px as 4P words][rows_ld rows] of plane ((n*C8+chunk)*2+part)*G+g;
          // out-of-range rows/columns are zero-filled and count towards the transaction bytes
          if (pw == 0) {
            if (lane == 0) {
              mbar_arrive_expect_tx(full, (uint32_t)(a.G * 4 * a.rows_ld) * row_bytes + b_stage_bytes);
              bulk_g2s(sa + a_stage_bytes, a.wpack + (size_t)k * (b_stage_bytes / 2), b_stage_bytes, full);
            }
            __syncwarp();
            if (lane < a.G * 4) {
              const int j = lane & 1, part = (lane >> 1) & 1, g = lane >> 2;
              const int pidx = ((n * a.c8[si] + (lone ? 0 : c0 + j)) * 2 + part) * a.G + g;
              const CUtensorMap* tm = si == 0 ? &tm0 : &tm1;
              asm volatile(
                  "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                  ::"r"(sa + lane * a_plane_bytes), "l"(reinterpret_cast<uint64_t>(tm)), "r"(4 * x0), "r"(y0), "r"(pidx), "r"(full)
                  : "memory");
            }
            __syncwarp();
          }
          if (++stage == (uint32_t)a.stages) { stage = 0; phase ^= 1; }
          continue;
        }
        if (pw == 0 && lane == 0) {
          mbar_arrive_expect_tx(full, (uint32_t)ncopy * row_bytes + b_stage_bytes);
          bulk_g2s(sa + a_stage_bytes, a.wpack + (size_t)k * (b_stage_bytes / 2), b_stage_bytes, full);
        }
        __syncwarp();
        for (int i = lane + 32 * pw; i < ncopy; i += 64) {
          const int pl = i / nrow, r = i - pl * nrow;          // pl = (g*2 + part)*2 + j
          const int j = pl & 1, part = (pl >> 1) & 1, g = pl >> 2;
          const int chunk = lone ? 0 : c0 + j;
          const __nv_bfloat16* gp = a.src[si] + (((size_t)n * a.c8[si] + chunk) * 2 + part) * a.slab_elems +
                                    (size_t)g * a.grp_elems + ((size_t)(y0 + r) * a.src_pitch + x0) * 8;
          bulk_g2s(sa + pl * a_plane_bytes + r * row_bytes, gp, row_bytes, full);
        }
        __syncwarp();
        if (++stage == (uint32_t)a.stages) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    constexpr uint32_t IDESC = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((128u >> 4) << 24);
    constexpr uint32_t IDESC2 = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)((2 * N) >> 3) << 17) | ((128u >> 4) << 24);
    // Warp-convergent issue: every lane runs the loop on warp-uniform values and one lane is elected inside
    // umma_bf16_ws, so descriptors stay in uniform registers and UTCHMMAs issue back to back.  (Under `if (lane == 0)`
    // every MMA cost an R2UR + ELECT + branch sequence, ~100 cycles of the issuing warp: more than an N <= 64 MMA.)
    uint32_t stage = 0, phase = 0, it = 0;
    constexpr uint32_t DESC_HI = (128u >> 4) | (1u << 14);       // SBO = 128 B, descriptor version 1
    constexpr uint32_t B_LBO = (uint32_t)(2 * N * 16 >> 4) << 16, BLO_OFF = (N * 16) >> 4;   // [j][hi|lo][N][16 B]
    const uint32_t a_lbo = (a_plane_bytes >> 4) << 16, alo_off = (2 * a_plane_bytes) >> 4;
    for (int tile = blockIdx.x; tile < a.ntiles; tile += gridDim.x, ++it) {
      const uint32_t buf = it & 1, tphase = (it >> 1) & 1;
      mbar_wait_warp(smem_u32(&s_tempty[buf]), tphase ^ 1);
      tc_fence_after();
      for (int k = 0; k < nk; ++k) {
        mbar_wait_warp(smem_u32(&s_full[stage]), phase);
        tc_fence_after();
        const uint32_t sa16 = (smem_base + stage * stage_bytes) >> 4;
        const uint32_t sb16 = sa16 + (a_stage_bytes >> 4);
        for (int t = 0; t < a.T; ++t) {
          const uint32_t dbase = tmem_base + buf * 256 + t * NACC;
          const uint32_t at16 = sa16 + (uint32_t)t * 128;
#pragma unroll
          for (int i = 0; i < TapTable<MODE>::NTAPS; ++i) {
            if (i >= a.ntaps) break;                                   // S2 without the 1x1 branch
            const uint32_t ao = TapTable<MODE>::a_off16(i, a.P, a_plane_bytes >> 4);
            const uint32_t d = dbase + TapTable<MODE>::dgrp(i) * a.T * NACC;
            const uint64_t a_hi = ((uint64_t)DESC_HI << 32) | ((at16 + ao) | a_lbo);
            const uint64_t a_lo = ((uint64_t)DESC_HI << 32) | ((at16 + ao + alo_off) | a_lbo);
            const uint32_t bo = sb16 + TapTable<MODE>::w_idx(i) * (B_TAP_BYTES >> 4);
            const uint64_t b_hi = ((uint64_t)DESC_HI << 32) | (bo | B_LBO);
            const uint64_t b_lo = ((uint64_t)DESC_HI << 32) | ((bo + BLO_OFF) | B_LBO);
            const uint32_t acc = (k > 0 || !TapTable<MODE>::first(i)) ? 1u : 0u;
            if (a.dbg & 2) continue;
            if (STACK) {
              umma_bf16_ws(d, a_hi, b_hi, IDESC2, acc);          // [hi*hi | hi*lo] : b_hi's 2N rows = [W_hi ; W_lo]
              umma_bf16_ws(d, a_lo, b_hi, IDESC, 1u);            // + lo*hi into the first half
            } else {
              umma_bf16_ws(d, a_hi, b_hi, IDESC, acc);
              umma_bf16_ws(d, a_lo, b_hi, IDESC, 1u);
              umma_bf16_ws(d, a_hi, b_lo, IDESC, 1u);
            }
          }
        }
        umma_commit_ws(smem_u32(&s_empty[stage]));                       // frees the smem slot when the MMAs retire
        if (k == nk - 1) umma_commit_ws(smem_u32(&s_tfull[buf]));       // accumulators of this tile complete
        if (++stage == (uint32_t)a.stages) { stage = 0; phase ^= 1; }
      }
    }
  } else {
    // ===================== epilogue (warps 2..5 = set 0, 7..10 = set 1) =====================
    const int q = warp & 3;                                    // TMEM lane quarter this warp may access
    const int eset = warp >= 7 ? 1 : 0;
    const int esplit = (TcEsets<N>::value == 1 || (a.dbg & 64)) ? 0 : 1;   // 0: everything goes to set 0
    uint32_t it = 0;
    const size_t oplane = (size_t)(a.OH + 2) * (a.OW + 2) * 8;
    const P8QGeom qg = P8QGeom::make(a.B, N, a.OH, a.OW);
    constexpr int nphase = (MODE == TC_T) ? 4 : 1;
    for (int tile = blockIdx.x; tile < a.ntiles; tile += gridDim.x, ++it) {
      const uint32_t buf = it & 1, tphase = (it >> 1) & 1;
      const int tx = tile % a.tiles_x, ty = (tile / a.tiles_x) % a.tiles_y, n = tile / (a.tiles_x * a.tiles_y);
      mbar_wait(smem_u32(&s_tfull[buf]), tphase);
      tc_fence_after();
      for (int t = 0; t < a.T; ++t) {
        const int m = t * 128 + q * 32 + lane;                 // flat staged position of this thread's anchor
        const int r = (int)__umulhi((uint32_t)m, a.P_magic), c = m - r * a.P;
        const int ay = ty * a.R + r, ax = tx * a.Wt + c;
        const bool valid = (r < a.R) && (c < a.Wt) && (ay < a.AH) && (ax < a.AW);
        for (int ph = 0; ph < nphase; ++ph) {
          if ((N / 32) % 2 == 1 && ((((t * nphase + ph) & 1) & esplit) != eset)) continue;   // odd block count: whole (t, ph) items alternate
          const int oy = (MODE == TC_T) ? 2 * ay + (ph >> 1) : ay;
          const int ox = (MODE == TC_T) ? 2 * ax + (ph & 1) : ax;
          __nv_bfloat16* obase = a.out + ((size_t)n * (N / 8) * 2) * oplane + ((size_t)(oy + 1) * (a.OW + 2) + (ox + 1)) * 8;
          const uint32_t tcol = tmem_base + ((uint32_t)(q * 32) << 16) + buf * 256;
#pragma unroll
          for (int cb = 0; cb < N / 32; ++cb) {
            if ((N / 32) % 2 == 0 && (((cb & 1) & esplit) != eset)) continue;           // even block count: the blocks alternate
            uint32_t v[32], w[32];
            tmem_ld32(tcol + (ph * a.T + t) * NACC + cb * 32, v);
            if (MODE == TC_S2 && a.ngrp == 2) tmem_ld32(tcol + (a.T + t) * NACC + cb * 32, w);   // S2 residual branch
            if (STACK) tmem_ld32(tcol + (ph * a.T + t) * NACC + N + cb * 32, w);                   // the hi*lo half
            tmem_ld_wait();
            if (STACK) {
#pragma unroll
              for (int e = 0; e < 32; ++e) v[e] = __float_as_uint(__uint_as_float(v[e]) + __uint_as_float(w[e]));
            }
            if (MODE == TC_S1 && a.out32) {
              // fp32 NCHW epilogue (the training data gradient, ops.py::_ConvTrain.backward): a lane owns one pixel, so every
              // channel's store is a 128-byte run across the warp; no P8 round trip through HBM
              if (valid && !(a.dbg & 1)) {
                size_t cs = (size_t)a.OH * a.OW;
                float* ob = a.out32 + ((size_t)n * a.o32_ctot + a.o32_c0 + cb * 32) * cs + (size_t)oy * a.OW + ox;
                if (a.o32_phase_c > 0) {
                  // phase-major output channels -> the interleaved full-resolution tensor (data gradient of a stride-2 conv)
                  const int chan = a.o32_c0 + cb * 32, ph = chan / a.o32_phase_c, cbase = chan - ph * a.o32_phase_c;
                  cs *= 4;
                  ob = a.out32 + ((size_t)n * a.o32_phase_c + cbase) * cs + (size_t)(2 * oy + (ph >> 1)) * (2 * a.OW) + 2 * ox + (ph & 1);
                }
#pragma unroll
                for (int e = 0; e < 32; ++e) {
                  const int o = cb * 32 + e;
                  float x = fmaf(__uint_as_float(v[e]), s_scale[o], s_shift[o]);
                  if (a.relu) x = fmaxf(x, 0.f);
                  ob[e * cs] = x;
                }
              }
            } else if (valid && !(a.dbg & 1)) {
#pragma unroll
              for (int ch = 0; ch < 4; ++ch) {
                float f[8];
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                  const int o = cb * 32 + ch * 8 + e;
                  float x = fmaf(__uint_as_float(v[ch * 8 + e]), s_scale[o], s_shift[o]);
                  if (a.relu) x = fmaxf(x, 0.f);
                  if (MODE == TC_S2 && a.ngrp == 2) x += __uint_as_float(w[ch * 8 + e]);
                  f[e] = x;
                }
                uint4 hi, lo;
                split8(f, hi, lo);
                __nv_bfloat16* p = obase + (size_t)(cb * 4 + ch) * 2 * oplane;
                *reinterpret_cast<uint4*>(p) = hi;
                *reinterpret_cast<uint4*>(p + oplane) = lo;
                if (a.outq) {
                  __nv_bfloat16* pq = a.outq + qg.slab(n, cb * 4 + ch, 0) + qg.pixel(oy, ox);
                  *reinterpret_cast<uint4*>(pq) = hi;
                  *reinterpret_cast<uint4*>(pq + qg.slab_elems()) = lo;
                }
              }
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

}  // namespace nconv

namespace nconv {

// ---- weight packing -> [kstage][wtap][j][part][CO][8] split bf16 -----------------------------------
//   S1: w [CO][CI][3][3]                     9 taps
//   S2: w [CO][CI][3][3] (+ w1 [CO][CI] 1x1) 9 (+1) taps
//   T : w [CI][CO][4][4]                     16 taps; `lead` leading input channels (the depth channel of
//       cat[depth, features], step2.py:174) take a K-stage of their own, zero-filled to 16.
__global__ void tc_pack_weights_kernel(const float* __restrict__ w, const float* __restrict__ w1,
                                       __nv_bfloat16* __restrict__ wp, int CO, int CI, int mode, int lead,
                                       int nstage, int ntw) {
  const int total = nstage * ntw * 2 * CO * 8;   // (kstage, wtap, j, co, e); both parts written together
  for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += gridDim.x * blockDim.x) {
    int e = idx & 7;
    int r = idx >> 3;
    int co = r % CO; r /= CO;
    int j = r & 1; r >>= 1;
    int wt = r % ntw;
    int ks = r / ntw;
    int ci;
    if (lead && ks < lead) ci = (j == 0 && e == 0) ? ks : -1;
    else ci = lead + (ks - lead) * 16 + j * 8 + e;
    float v = 0.f;
    if (ci >= 0 && ci < CI) {
      if (mode == TC_T) v = __ldg(w + ((size_t)ci * CO + co) * 16 + wt);
      else if (wt < 9) v = __ldg(w + ((size_t)co * CI + ci) * 9 + wt);
      else v = __ldg(w1 + (size_t)co * CI + ci);
    }
    __nv_bfloat16 hi, lo;
    split_bf16(v, hi, lo);
    const size_t base = ((size_t)(ks * ntw + wt) * 2) * 2 * CO * 8;     // [j][part][CO][8]
    wp[base + ((size_t)(j * 2 + 0) * CO + co) * 8 + e] = hi;
    wp[base + ((size_t)(j * 2 + 1) * CO + co) * 8 + e] = lo;
  }
}

static inline int tc_lead(int mode, int CA) { return (mode == TC_T && CA == 1) ? 1 : 0; }
static inline int tc_nstage(int mode, int CA, int CB) { return tc_lead(mode, CA) ? 1 + CB / 16 : (CA + CB) / 16; }
static inline int tc_ntaps_w(int mode, bool has_1x1) { return mode == TC_T ? 16 : (mode == TC_S2 && has_1x1 ? 10 : 9); }

// ---- tile geometry ---------------------------------------------------------------------------------
struct TcPlan { int R, Wt, P, T, rows_alloc, rows_ld, G, ngrp, tiles_x, tiles_y; size_t smem; bool ok; int stages; double cost; };

static TcPlan tc_plan_stages(int mode, int AH, int AW, int N, int ngrp, int ntw, bool stack, int stages) {
  TcPlan best = {}; best.ok = false;
  double best_cost = 1e300;
  const int tmax = 256 / (N * (stack ? 2 : 1) * ngrp);  // M-tiles per TMEM buffer (256 columns each)
  const int hc = (mode == TC_S2) ? 1 : 2, hr = hc;     // halo columns / rows of the staged tile
  const int G = (mode == TC_S2) ? 4 : 1;
  for (int Wt = 8; Wt <= 250 && Wt <= AW + 7; ++Wt) {
    const int P = Wt + hc;
    for (int R = 1; R <= 40; ++R) {
      const int flat = R * P - hc;
      const int T = (flat + 127) / 128;
      if (T > tmax) break;
      const int maxshift = hr * P + hc;
      int rows_alloc = (T * 128 + maxshift + P) / P;
      if (mode == TC_S2) while ((rows_alloc * P) % 8) ++rows_alloc;   // 128 B aligned planes (TMA tensor copies)
      const size_t smem = (size_t)stages * ((size_t)G * 4 * rows_alloc * P * 16 + (size_t)ntw * 2 * 2 * N * 16);
      if (smem > (size_t)TC_SMEM_MAX) continue;
      const int tx = (AW + Wt - 1) / Wt, ty = (AH + R - 1) / R;
      // MMA rows issued per tile (x taps, constant) plus a load term for re-read halo rows and weights
      const double cost = (double)tx * ty * (T * 128.0 + 0.25 * G * (R + hr) * P + 0.02 * ntw * N);
      if (cost < best_cost) {
        best_cost = cost;
        best = {R, Wt, P, T, rows_alloc, R + hr, G, ngrp, tx, ty, smem, true, stages, cost};
      }
    }
  }
  return best;
}

// A 3-deep ring lets the loads run two K-stages ahead of the MMAs; take it unless it forces a much worse tile.
static TcPlan tc_plan(int mode, int AH, int AW, int N, int ngrp, int ntw, bool stack) {
  static const char* force = getenv("NCONV_B200_TC_STAGES");          // A/B measurements only
  const TcPlan p2 = tc_plan_stages(mode, AH, AW, N, ngrp, ntw, stack, 2);
  const TcPlan p3 = tc_plan_stages(mode, AH, AW, N, ngrp, ntw, stack, 3);
  if (force) return atoi(force) == 3 && p3.ok ? p3 : p2;
  if (p3.ok && (!p2.ok || p3.cost <= 1.12 * p2.cost)) return p3;
  return p2;
}

template <int N, int MODE, bool STACK = false>
static int launch_tc(const TcArgs& a, const TcPlan& p, cudaStream_t st) {
  auto kern = conv_tc_kernel<N, MODE, STACK>;
  static bool attr_done[NCONV_MAX_DEVICES] = {};
  static int sm_count[NCONV_MAX_DEVICES] = {};
  int dev = 0, sms = 0;
  NCONV_RETURN_IF_ERR(cudaGetDevice(&dev));
  if (dev >= NCONV_MAX_DEVICES || !attr_done[dev]) {
    NCONV_RETURN_IF_ERR(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM_MAX));
    if (dev < NCONV_MAX_DEVICES) attr_done[dev] = true;
  }
  if (dev < NCONV_MAX_DEVICES) sms = sm_count[dev];
  if (sms == 0) {
    NCONV_RETURN_IF_ERR(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    if (dev < NCONV_MAX_DEVICES) sm_count[dev] = sms;
  }
  const int grid = a.ntiles < sms ? a.ntiles : sms;
  const int nk = a.pairs[0] + a.pairs[1];
  const double opx = (double)a.B * a.OH * a.OW, apx = (double)a.B * a.AH * a.AW;
  const double ipx = (a.mode == TC_S2) ? 4.0 * apx : apx;
  const int CI = 16 * nk;
  // flops: real taps per output pixel (9, 9+1, or 4 of the 16 per output phase)
  const double taps_per_out = (a.mode == TC_T) ? 4.0 : (double)a.ntaps;
  alignas(64) CUtensorMap tm[2];
  memset(tm, 0, sizeof(tm));
  if (a.use_tmap) {
    for (int si = 0; si < 2; ++si) {
      const int use = (a.src[si] && a.c8[si] > 0) ? si : 0;
      const int rc = tc_make_tmap(&tm[si], a.src[use], a.B * a.c8[use] * 2 * a.G, a.src_rows, a.src_pitch, a.P, a.rows_ld);
      if (rc) return rc;
    }
  }
  char label[64];
  snprintf(label, sizeof(label), "tc_conv%s %d->%d @%dx%d", a.mode == TC_S1 ? "3x3" : (a.mode == TC_S2 ? "3x3s2" : "T4x4s2"),
           CI, N, a.OH, a.OW);
  {
    // algorithmic bytes: fp32-equivalent activations in and out (P8 holds the same 4 B per value)
    LaunchScope ls(st, label, 2.0 * CI * N * taps_per_out * opx, 4.0 * (CI * ipx + N * opx * (a.outq ? 2.0 : 1.0)));
    kern<<<grid, 224 + 128 * (TcEsets<N>::value - 1), p.smem, st>>>(a, tm[0], tm[1]);
  }
  return (int)cudaGetLastError();
}

static void tc_build_taps(TcArgs& a, int mode, bool has_1x1) {
  const uint32_t plane = (uint32_t)a.rows_alloc * a.P * 16;
  int n = 0;
  if (mode == TC_S1) {
    for (int u = 0; u < 3; ++u)
      for (int v = 0; v < 3; ++v) {
        const TcTap tp = {(uint32_t)(u * a.P + v) * 16, (uint16_t)(u * 3 + v), 0, (uint8_t)(n == 0)};
        a.taps[n++] = tp;
      }
  } else if (mode == TC_S2) {
    // input row 2y+u-1: u=0 -> odd phase, phase row y-1 (staged row r); u=1 -> even phase, row y
    // (staged r+1); u=2 -> odd phase, row y (staged r+1).  Same along x.
    for (int u = 0; u < 3; ++u)
      for (int v = 0; v < 3; ++v) {
        const int py = (u == 1) ? 0 : 1, rs = (u == 0) ? 0 : 1;
        const int px = (v == 1) ? 0 : 1, cs = (v == 0) ? 0 : 1;
        const TcTap tp = {(uint32_t)(py * 2 + px) * 4 * plane + (uint32_t)(rs * a.P + cs) * 16, (uint16_t)(u * 3 + v), 0,
                          (uint8_t)(n == 0)};
        a.taps[n++] = tp;
      }
    if (has_1x1) {
      const TcTap tp = {(uint32_t)(a.P + 1) * 16, 9, 1, 1};                // x[2y, 2x]: phase (0,0)
      a.taps[n++] = tp;
    }
  } else {
    // out[2a+py, 2b+px] += in[a+py-t, b+px-s] * W[.., (1-py)+2t, (1-px)+2s]
    bool seen[4] = {false, false, false, false};
    for (int kh = 0; kh < 4; ++kh)
      for (int kw = 0; kw < 4; ++kw) {
        const int py = 1 - (kh & 1), dy = py - (kh >> 1);
        const int px = 1 - (kw & 1), dx = px - (kw >> 1);
        const int g = py * 2 + px;
        const TcTap tp = {(uint32_t)((1 + dy) * a.P + (1 + dx)) * 16, (uint16_t)(kh * 4 + kw), (uint8_t)g, (uint8_t)(!seen[g])};
        a.taps[n++] = tp;
        seen[g] = true;
      }
  }
  a.ntaps = n;
}

}  // namespace nconv

#include "tc_conv_x.cuh"

using namespace nconv;

// Co = 32 stride-1 layers with >= 64 input channels run on the pixels-as-N kernel (tc_conv_x.cuh) with its own weight
// packing.  With 32 input channels the pixels-on-M kernel (N-stacked [W_hi | W_lo]) is the faster one since round 2's
// convergent MMA issue (32->32 @480x640: 0.68 vs 0.79 ms; 64->32: 1.21 vs 1.20).  NCONV_B200_NO_TCX=1 keeps every layer on
// conv_tc_kernel, NCONV_B200_TCX_ALL=1 sends every Co = 32 layer to conv_tcx (A/B measurements only; read once per
// process, so packing and launch always agree).
static bool tcx_on(int mode, int CO, int CI) {
  static const bool off = getenv("NCONV_B200_NO_TCX") != nullptr;
  static const bool all = getenv("NCONV_B200_TCX_ALL") != nullptr;
  // CI <= 64: the kernel keeps ALL packed weights resident in shared memory (24 KB per 16 input channels); wider Co = 32
  // layers (none in the reference's stack) take the pixels-on-M kernel, which streams its weights per K-stage
  return !off && mode == TC_S1 && CO == TCX_CO && CI <= 64 && (all || CI >= 64);
}

extern "C" size_t nconv_b200_tc_wpack_bytes(int mode, int CO, int CA, int CB, int has_1x1) {
  if (CO <= 0 || CA <= 0 || CB < 0 || mode < 0 || mode > 2) return 0;
  if (tcx_on(mode, CO, CA + CB)) return (size_t)((CA + CB) / 16) * TCX_WSTAGE_BYTES;
  return (size_t)tc_nstage(mode, CA, CB) * tc_ntaps_w(mode, has_1x1 != 0) * 2 * 2 * CO * 8 * sizeof(__nv_bfloat16);
}

extern "C" int nconv_b200_tc_pack_weights(int mode, const float* w, const float* w1, void* wpack, int CO, int CA,
                                          int CB, void* stream) {
  if (!w || !wpack || CO <= 0 || CA <= 0 || CB < 0 || mode < 0 || mode > 2) return NCONV_ERR_BAD_ARG;
  const int lead = tc_lead(mode, CA);
  if ((!lead && CA % 16) || CB % 16) return NCONV_ERR_UNSUPPORTED;
  if (w1 && mode != TC_S2) return NCONV_ERR_BAD_ARG;
  if (tcx_on(mode, CO, CA + CB)) {
    const int nst = (CA + CB) / 16, tot = nst * 3 * 4 * 128 * 8;
    {
      LaunchScope ls((cudaStream_t)stream, "tc_pack_weights", 0.0, 8.0 * CO * (CA + CB) * 9);
      tcx_pack_weights_kernel<<<(tot + 255) / 256, 256, 0, (cudaStream_t)stream>>>(w, (__nv_bfloat16*)wpack, CA + CB, nst);
    }
    return (int)cudaGetLastError();
  }
  const int nstage = tc_nstage(mode, CA, CB), ntw = tc_ntaps_w(mode, w1 != nullptr);
  const int total = nstage * ntw * 2 * CO * 8;
  {
    LaunchScope ls((cudaStream_t)stream, "tc_pack_weights", 0.0, 8.0 * CO * (CA + CB) * ntw);
    tc_pack_weights_kernel<<<(total + 255) / 256, 256, 0, (cudaStream_t)stream>>>(w, w1, (__nv_bfloat16*)wpack, CO,
                                                                                 CA + CB, mode, lead, nstage, ntw);
  }
  return (int)cudaGetLastError();
}

extern "C" int nconv_b200_conv_tc(const nconv_b200_tc_desc* d, void* stream) {
  if (!d || !d->a || !d->wpack || (!d->y && !d->y_f32)) return NCONV_ERR_BAD_ARG;
  if (d->B <= 0 || d->H <= 0 || d->W <= 0 || d->CA <= 0 || d->CB < 0) return NCONV_ERR_BAD_ARG;
  if (d->mode < 0 || d->mode > 2) return NCONV_ERR_BAD_ARG;
  if (d->y_f32) {
    if (d->mode != TC_S1 || d->yq || d->y_c0 < 0) return NCONV_ERR_BAD_ARG;
    if (d->y_phase_c > 0 ? ((d->y_phase_c % 32) || (d->y_c0 % 32) || d->y_c0 + d->CO > 4 * d->y_phase_c) : (d->y_c0 + d->CO > d->y_ctot))
      return NCONV_ERR_BAD_ARG;
    if (tcx_on(d->mode, d->CO, d->CA + d->CB)) return NCONV_ERR_UNSUPPORTED;       // the pixels-on-N kernel writes P8 only
  }
  const int mode = d->mode;
  const int lead = tc_lead(mode, d->CA);
  if ((!lead && d->CA % 16) || d->CB % 16 || (d->CB > 0 && !d->b)) return NCONV_ERR_UNSUPPORTED;
  if (!(d->CO == 32 || d->CO == 64)) return NCONV_ERR_UNSUPPORTED;
  if (mode == TC_S2 && ((d->H | d->W) & 1)) return NCONV_ERR_UNSUPPORTED;
  if (d->yq && (mode == TC_T ? 0 : ((mode == TC_S2 ? (d->H / 2) | (d->W / 2) : (d->H | d->W)) & 1))) return NCONV_ERR_UNSUPPORTED;
  const bool has_1x1 = (mode == TC_S2) && d->has_1x1;
  const int AH = (mode == TC_S2) ? d->H / 2 : d->H, AW = (mode == TC_S2) ? d->W / 2 : d->W;
  if (tcx_on(mode, d->CO, d->CA + d->CB)) {
    const int nk = (d->CA + d->CB) / 16;
    const TcxPlan p = tcx_plan(AH, AW, nk);
    if (!p.ok) return NCONV_ERR_UNSUPPORTED;
    TcxArgs a = {};
    a.src[0] = (const __nv_bfloat16*)d->a; a.src[1] = (const __nv_bfloat16*)d->b;
    a.c8[0] = d->CA / 8; a.c8[1] = d->CB / 8;
    a.pairs[0] = d->CA / 16; a.pairs[1] = d->CB / 16;
    a.slab_elems = (size_t)(d->H + 2) * (d->W + 2) * 8; a.src_pitch = d->W + 2; a.src_rows = d->H + 2;
    a.wpack = (const __nv_bfloat16*)d->wpack; a.scale = d->scale; a.shift = d->shift;
    a.out = (__nv_bfloat16*)d->y; a.outq = (__nv_bfloat16*)d->yq;
    a.B = d->B; a.AH = AH; a.AW = AW;
    a.R = p.R; a.Wt = p.Wt; a.P = p.P; a.rows_alloc = p.rows_alloc; a.rows_ld = p.rows_ld;
    a.tiles_x = p.tiles_x; a.tiles_y = p.tiles_y; a.ntiles = p.tiles_x * p.tiles_y * d->B;
    a.relu = d->relu; a.stages = p.stages;
    { static const char* dbg = getenv("NCONV_B200_TC_DEBUG"); a.dbg = dbg ? atoi(dbg) : 0; }
    return launch_tcx(a, p, (cudaStream_t)stream);
  }
  const int ngrp = (mode == TC_T) ? 4 : (has_1x1 ? 2 : 1);
  const int ntw = tc_ntaps_w(mode, has_1x1);
  static const bool no_stack = getenv("NCONV_B200_NO_STACK") != nullptr;      // A/B measurements only
  const bool stack = (mode == TC_S1) && d->CO == 32 && !no_stack;
  const TcPlan p = tc_plan(mode, AH, AW, d->CO, ngrp, ntw, stack);
  if (!p.ok) return NCONV_ERR_UNSUPPORTED;
  TcArgs a = {};
  a.src[0] = (const __nv_bfloat16*)d->a; a.src[1] = (const __nv_bfloat16*)d->b;
  a.c8[0] = lead ? 1 : d->CA / 8; a.c8[1] = d->CB / 8;
  a.pairs[0] = lead ? 1 : d->CA / 16; a.pairs[1] = d->CB / 16;
  if (mode == TC_S2) {
    const P8QGeom q = P8QGeom::make(d->B, 8, d->H, d->W);
    a.slab_elems = q.slab_elems(); a.grp_elems = q.phase_elems(); a.src_pitch = q.Wq1; a.src_rows = q.Hq1;
  } else {
    a.slab_elems = (size_t)(d->H + 2) * (d->W + 2) * 8; a.grp_elems = 0; a.src_pitch = d->W + 2; a.src_rows = d->H + 2;
  }
  a.wpack = (const __nv_bfloat16*)d->wpack; a.scale = d->scale; a.shift = d->shift;
  a.out = (__nv_bfloat16*)d->y; a.outq = (__nv_bfloat16*)d->yq;
  a.out32 = d->y_f32; a.o32_ctot = d->y_ctot; a.o32_c0 = d->y_c0; a.o32_phase_c = d->y_f32 ? d->y_phase_c : 0;
  a.B = d->B; a.AH = AH; a.AW = AW;
  a.OH = (mode == TC_T) ? 2 * d->H : AH; a.OW = (mode == TC_T) ? 2 * d->W : AW;
  a.R = p.R; a.Wt = p.Wt; a.P = p.P; a.T = p.T; a.rows_alloc = p.rows_alloc; a.rows_ld = p.rows_ld;
  a.P_magic = (uint32_t)(((1ull << 32) + (uint64_t)p.P - 1) / (uint64_t)p.P);
  a.G = p.G; a.ngrp = ngrp; a.ntaps_w = ntw;
  a.tiles_x = p.tiles_x; a.tiles_y = p.tiles_y; a.ntiles = p.tiles_x * p.tiles_y * d->B;
  a.relu = d->relu; a.mode = mode; a.stages = p.stages;
  { static const char* dbg = getenv("NCONV_B200_TC_DEBUG"); a.dbg = dbg ? atoi(dbg) : 0; }
  {
    // Tensor-map loads pay where a stage is many short rows: the stride-2 layers stage 16 planes of 17-pixel rows
    // (32->64 s2 @240x320: 0.71 -> 0.50 ms).  The stride-1 / transposed layers measured 0-9 % slower with them
    // (one box completes as a whole; row copies trickle in), so they keep per-row bulk copies.
    // NCONV_B200_TC_TMAP=0/1 forces either path (A/B measurements only).
    static const char* force = getenv("NCONV_B200_TC_TMAP");
    a.use_tmap = force ? (atoi(force) != 0) : (mode == TC_S2);
    if (p.P * 4 > 256 || (p.rows_alloc * p.P) % 8) a.use_tmap = 0;
  }
  tc_build_taps(a, mode, has_1x1);
  cudaStream_t st = (cudaStream_t)stream;
  if (mode == TC_S1) {
    if (d->CO == 64) return launch_tc<64, TC_S1>(a, p, st);
    return stack ? launch_tc<32, TC_S1, true>(a, p, st) : launch_tc<32, TC_S1>(a, p, st);
  }
  if (mode == TC_S2) return d->CO == 64 ? launch_tc<64, TC_S2>(a, p, st) : launch_tc<32, TC_S2>(a, p, st);
  return d->CO == 64 ? launch_tc<64, TC_T>(a, p, st) : launch_tc<32, TC_T>(a, p, st);
}
