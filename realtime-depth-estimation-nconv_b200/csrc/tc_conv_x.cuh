// tcgen05 implicit-GEMM 3x3/s1 convolution for Co = 32 in the "pixels as N" orientation
// (ConvBlock / NewFusionBlock / Basic2d with 32 output channels, /root/reference/models/step2.py:
// 178-195, 216-236, 290-297).  Included by tc_conv.cu.
//
// Why a second orientation.  conv_tc_kernel puts 128 pixels on M and Co on N; at Co = 32 an MMA is
// M128 x N64 x K16 and is bound by reading its 4 KB pixel tile from shared memory (~80 cycles against
// a 32-cycle tensor floor).  Here the roles are swapped:
//
//   D[128 rows][256 pixels] += A[128 rows][K16] (weights, shared memory) * B[256 pixels][K16]^T
//
// * B is the staged halo tile exactly as before (P8 planes, flat with pitch P; a tap is a start
//   address shift), now 256 consecutive staged pixels per MMA.
// * The 128 weight rows are the three horizontal taps of one kernel row stacked:
//   row = g*32 + co  ->  output channel co, tap (dy, dx = g); rows 96..127 are zero (never read back).
//   All three taps multiply the SAME pixel columns, so tap g's result for output pixel m sits in
//   column m+g: the epilogue adds three column-shifted rows (a 256-column tile yields 254 outputs).
// * bf16x3 by K-stacking: the K=16 slice of the "hi" MMAs is [x_hi(8 ch) ; x_lo(8 ch)] (descriptor
//   LBO = hi->lo plane distance) against [W_hi ; W_hi], i.e. W_hi*(x_hi+x_lo) in ONE MMA per
//   8-channel chunk; a third MMA per 16 channels is W_lo * x_hi.  Per kernel row and 16 input
//   channels: 3 MMAs of 128x256x16 = 384 cycles for 256 pixels, against 3 taps x 2 MMAs x ~80
//   cycles x 2 (128-pixel tiles) = 960 cycles in the pixel-M kernel.
// * All packed weights of the layer stay resident in shared memory (24 KB per 16 input channels: the
//   [W_hi ; W_hi] tiles alias their two K-halves with LBO = 0); only pixel stages stream through the ring,
//   one TMA tensor copy per staged plane.
//
// Warp roles (16 warps): warp 0 producer (weights once, then 4 tensor copies per K-stage), warp 1 MMA issuer, six
// STAGER warps in two sets (2, 4, 5 and 6, 8, 9: TMEM lane quarters 2, 0, 1 = the three tap groups; a warp may only read
// quarter warp % 4; the sets take alternate windows so one set's shared-memory stores overlap the other's tcgen05.ld:
// draining 3 x 32 KB per tile at the 16 B/clk a lane quarter delivers is the epilogue's floor) and eight EMITTER warps
// (3, 7, 10..15) in two groups of four.  Stagers drain the accumulator window by window (32 columns)
// into a 128-row shared-memory ring (one plane per tap group, row = column - tap shift); emitters sum the three planes,
// apply scale/shift/ReLU, split to hi/lo and store P8 (warp -> 8-channel chunk, lane -> pixel), the two groups taking
// alternate windows.  Stagers and emitters are decoupled by per-window mbarriers (full: 3 stager arrivals, empty: 4
// emitter arrivals), three windows may be in flight; round 1's version ran both halves in the same four warps with four
// 128-thread barriers per two windows and was latency-bound (5 200 cycles of epilogue per tile against 3 000 of MMAs).
// TMEM: two 256-column accumulators (tile i's epilogue overlaps tile i+1's MMAs).
#pragma once
#include <cuda.h>

namespace nconv {

constexpr int TCX_THREADS = 512;                 // 16 warps
constexpr int TCX_N = 256;                 // pixels per MMA / TMEM accumulator columns
constexpr int TCX_OUT = TCX_N - 2;         // output pixels per tile (column shifts 0..2)
constexpr int TCX_WHALF_BYTES = 128 * 16;              // one K-half of an A tile: [128 rows][8] bf16
constexpr int TCX_WROW_BYTES = 4 * TCX_WHALF_BYTES;    // per kernel row: hi j0 | hi j1 | lo (2 K-halves)
constexpr int TCX_WSTAGE_BYTES = 3 * TCX_WROW_BYTES;   // per K-stage (16 input channels): 24 KB
constexpr int TCX_MAX_STAGES = 8;
constexpr int TCX_S_PITCH = 36;                        // floats per staged-output row (32 co + pad: conflict-free 128-bit reads)
constexpr int TCX_S_ROWS = 128;                        // ring rows = 4 windows
constexpr int TCX_S_PLANE = TCX_S_ROWS * TCX_S_PITCH;  // floats per tap-group plane
constexpr int TCX_S_BYTES = (3 * TCX_S_PLANE * 4 + 127) / 128 * 128;   // the ring behind it must stay 128 B aligned
constexpr int TCX_CO = 32;

struct TcxArgs {
  const __nv_bfloat16* src[2];
  int c8[2], pairs[2];
  size_t slab_elems;
  int src_pitch, src_rows;
  const __nv_bfloat16* wpack;   // [kstage][dy][hi j0 | hi j1 | lo k0 | lo k1][128][8]
  const float* scale; const float* shift;
  __nv_bfloat16* out; __nv_bfloat16* outq;
  int B, AH, AW;
  int R, Wt, P, rows_alloc, rows_ld;
  int tiles_x, tiles_y, ntiles;
  int relu, stages;
  int dbg;                      // NCONV_B200_TC_DEBUG bits (timing experiments only): 1 no epilogue work, 2 no MMAs, 4 no loads
};

__global__ void __launch_bounds__(TCX_THREADS, 1) conv_tcx_kernel(const __grid_constant__ TcxArgs a, const __grid_constant__ CUtensorMap tm0,
                                                                 const __grid_constant__ CUtensorMap tm1) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t s_full[TCX_MAX_STAGES], s_empty[TCX_MAX_STAGES], s_tfull[2], s_tempty[2], s_wfull;
  __shared__ __align__(8) uint64_t s_sfull[4], s_sempty[4];          // staging ring: window slot staged / emitted
  __shared__ uint32_t s_tmem_base;
  __shared__ float s_scale[TCX_CO], s_shift[TCX_CO];

  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;   // provably warp-uniform
  const int nk = a.pairs[0] + a.pairs[1];
  const uint32_t plane_bytes = (uint32_t)a.rows_alloc * a.P * 16;    // one (part, j) plane of a stage
  const uint32_t stage_bytes = 4 * plane_bytes;
  const uint32_t w_bytes = (uint32_t)nk * TCX_WSTAGE_BYTES;
  const uint32_t smem_base = smem_u32(smem);
  const uint32_t s_out = smem_base + w_bytes;                        // epilogue staging (3 planes)
  const uint32_t ring = s_out + TCX_S_BYTES;

  if (threadIdx.x < TCX_CO) {
    s_scale[threadIdx.x] = a.scale ? __ldg(a.scale + threadIdx.x) : 1.f;
    s_shift[threadIdx.x] = a.shift ? __ldg(a.shift + threadIdx.x) : 0.f;
  }
  if (threadIdx.x == 0) {
    for (int s = 0; s < TCX_MAX_STAGES; ++s) { mbar_init(smem_u32(&s_full[s]), 1); mbar_init(smem_u32(&s_empty[s]), 1); }
    for (int b = 0; b < 2; ++b) { mbar_init(smem_u32(&s_tfull[b]), 1); mbar_init(smem_u32(&s_tempty[b]), 6); }
    for (int b = 0; b < 4; ++b) { mbar_init(smem_u32(&s_sfull[b]), 3); mbar_init(smem_u32(&s_sempty[b]), 4); }
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

  if (warp == 0) {
    // ===================== producer =====================
    const int pw = 0;
    if (pw == 0 && lane == 0) {
      // the layer's packed weights, once per CTA
      const uint32_t wf = smem_u32(&s_wfull);
      mbar_arrive_expect_tx(wf, w_bytes);
      for (int k = 0; k < nk; ++k)
        bulk_g2s(smem_base + k * TCX_WSTAGE_BYTES, (const uint8_t*)a.wpack + (size_t)k * TCX_WSTAGE_BYTES, TCX_WSTAGE_BYTES, wf);
    }
    // One tensor copy per staged plane: box = [P px (as 4P 32-bit words, contiguous)][R+2 rows] of plane
    // ((n*C8 + chunk)*2 + part); the innermost box extent is a whole 16*P-byte row, not one 16-byte pixel.
    // (Per-row cp.async.bulk copies of < 1 KB ran at ~77 cycles each per SM whatever the number of issuing
    // warps - 2.5 TB/s over the chip - and were the kernel's bottleneck; out-of-bounds parts of a box are
    // zero-filled by the TMA unit and count towards the transaction bytes.)
    uint32_t stage = 0, phase = 0;
    const uint32_t box_bytes = (uint32_t)a.rows_ld * a.P * 16;
    if (pw == 0) {
      for (int tile = blockIdx.x; tile < a.ntiles; tile += gridDim.x) {
        const int tx = tile % a.tiles_x, ty = (tile / a.tiles_x) % a.tiles_y, n = tile / (a.tiles_x * a.tiles_y);
        const int y0 = ty * a.R, x0 = tx * a.Wt;
        for (int k = 0; k < nk; ++k) {
          const int si = (k < a.pairs[0]) ? 0 : 1;
          const int c0 = 2 * (si == 0 ? k : k - a.pairs[0]);
          mbar_wait(smem_u32(&s_empty[stage]), phase ^ 1);
          const uint32_t full = smem_u32(&s_full[stage]);
          const uint32_t sa = ring + stage * stage_bytes;
          if (a.dbg & 4) { if (lane == 0) mbar_arrive(full); __syncwarp(); if (++stage == (uint32_t)a.stages) { stage = 0; phase ^= 1; } continue; }
          if (lane == 0) mbar_arrive_expect_tx(full, 4 * box_bytes);
          __syncwarp();
          if (lane < 4) {
            const int j = lane & 1, part = lane >> 1;          // staged plane = part*2 + j
            const int pidx = ((n * a.c8[si] + c0 + j) * 2 + part);
            const CUtensorMap* tm = si == 0 ? &tm0 : &tm1;
            asm volatile(
                "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                ::"r"(sa + lane * plane_bytes), "l"(reinterpret_cast<uint64_t>(tm)), "r"(4 * x0), "r"(y0), "r"(pidx), "r"(full)
                : "memory");
          }
          __syncwarp();
          if (++stage == (uint32_t)a.stages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    constexpr uint32_t IDESC = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(TCX_N >> 3) << 17) | ((128u >> 4) << 24);
    constexpr uint32_t DESC_HI = (128u >> 4) | (1u << 14);          // SBO = 128 B, descriptor version 1
    constexpr uint32_t W_LBO = (uint32_t)(TCX_WHALF_BYTES >> 4) << 16;   // lo tile: K-halves 2 KB apart
    // hi tiles: LBO = 0, both K-halves read the SAME 2 KB [W_hi ; W_hi] (against [x_hi ; x_lo])
    // warp-convergent issue (see tc_conv.cu): uniform waits, one lane elected inside umma_bf16_ws
    mbar_wait_warp(smem_u32(&s_wfull), 0);
    tc_fence_after();
    uint32_t stage = 0, phase = 0, it = 0;
    const uint32_t x_lbo_hl = ((2 * plane_bytes) >> 4) << 16;       // hi plane -> lo plane of the same chunk
    const uint32_t x_lbo_jj = (plane_bytes >> 4) << 16;             // chunk j0 -> j1 (both hi)
    for (int tile = blockIdx.x; tile < a.ntiles; tile += gridDim.x, ++it) {
      const uint32_t buf = it & 1, tphase = (it >> 1) & 1;
      mbar_wait_warp(smem_u32(&s_tempty[buf]), tphase ^ 1);
      tc_fence_after();
      const uint32_t d = tmem_base + buf * TCX_N;
      for (int k = 0; k < nk; ++k) {
        mbar_wait_warp(smem_u32(&s_full[stage]), phase);
        tc_fence_after();
        const uint32_t x16 = (ring + stage * stage_bytes) >> 4;
        const uint32_t w16 = (smem_base + (uint32_t)k * TCX_WSTAGE_BYTES) >> 4;
        if (!(a.dbg & 2)) {
#pragma unroll
          for (int dy = 0; dy < 3; ++dy) {
            const uint32_t xs = x16 + (uint32_t)(dy * a.P);
            const uint32_t ws = w16 + (uint32_t)dy * (TCX_WROW_BYTES >> 4);
            const uint64_t x_h0 = ((uint64_t)DESC_HI << 32) | (xs | x_lbo_hl);                          // [x_hi ; x_lo] chunk j0
            const uint64_t x_h1 = ((uint64_t)DESC_HI << 32) | ((xs + (plane_bytes >> 4)) | x_lbo_hl);   // chunk j1
            const uint64_t x_16 = ((uint64_t)DESC_HI << 32) | (xs | x_lbo_jj);                          // x_hi, 16 channels
            const uint64_t w_h0 = ((uint64_t)DESC_HI << 32) | ws;
            const uint64_t w_h1 = ((uint64_t)DESC_HI << 32) | (ws + (TCX_WHALF_BYTES >> 4));
            const uint64_t w_lo = ((uint64_t)DESC_HI << 32) | ((ws + 2 * (TCX_WHALF_BYTES >> 4)) | W_LBO);
            umma_bf16_ws(d, w_h0, x_h0, IDESC, (k > 0 || dy > 0) ? 1u : 0u);
            umma_bf16_ws(d, w_h1, x_h1, IDESC, 1u);
            umma_bf16_ws(d, w_lo, x_16, IDESC, 1u);
          }
        }
        umma_commit_ws(smem_u32(&s_empty[stage]));
        if (k == nk - 1) umma_commit_ws(smem_u32(&s_tfull[buf]));
        if (++stage == (uint32_t)a.stages) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 2 || warp == 4 || warp == 5 || warp == 6 || warp == 8 || warp == 9) {
    // ===================== stagers: TMEM -> shared-memory ring =====================
    // TMEM lane 32g + co holds tap-group g of output channel co: the warp of lane quarter g copies its 32x32 block of
    // every 32-column window into plane g of the ring at row (column - g) (conflict-free 128-byte stores per column).
    const int q = warp & 3;                                          // TMEM lane quarter = tap group (0, 1, 2)
    const uint32_t sj = warp >= 6 ? 1u : 0u;                         // stager set: windows with (running index) % 2 == sj
    float* S_st = reinterpret_cast<float*>(smem + w_bytes) + q * TCX_S_PLANE + lane;
    uint32_t it = 0;                                                 // tile counter; running window index gw = 8 it + w
    for (int tile = blockIdx.x; tile < a.ntiles; tile += gridDim.x, ++it) {
      const uint32_t buf = it & 1, tphase = (it >> 1) & 1;
      const uint32_t tcol = tmem_base + ((uint32_t)(q * 32) << 16) + buf * TCX_N;
      mbar_wait(smem_u32(&s_tfull[buf]), tphase);
      tc_fence_after();
      uint32_t va[32], vb[32];
      auto stage_window = [&](const uint32_t* v, uint32_t gw) {
        // window gw overwrites the ring rows of window gw - 4, which the emitters of windows gw - 4 and gw - 3 read
        if (gw >= 4) mbar_wait(smem_u32(&s_sempty[(gw - 4) & 3]), ((gw - 4) >> 2) & 1);
        if (gw >= 3) mbar_wait(smem_u32(&s_sempty[(gw - 3) & 3]), ((gw - 3) >> 2) & 1);
        if (!(a.dbg & 8)) {
#pragma unroll
          for (int e = 0; e < 32; ++e) S_st[((32 * gw + e - q) & (TCX_S_ROWS - 1)) * TCX_S_PITCH] = __uint_as_float(v[e]);
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_u32(&s_sfull[gw & 3]));
      };
      if (!(a.dbg & 1)) {
        // this set's windows of the tile: w = sj, sj + 2, sj + 4, sj + 6 (the running index 8 it + w has w's parity)
        const uint32_t g0 = 8 * it + sj;
        tmem_ld32(tcol + sj * 32, va);
        tmem_ld_wait();
        tmem_ld32(tcol + (sj + 2) * 32, vb);
        stage_window(va, g0);
        tmem_ld_wait();
        tmem_ld32(tcol + (sj + 4) * 32, va);
        stage_window(vb, g0 + 2);
        tmem_ld_wait();
        tmem_ld32(tcol + (sj + 6) * 32, vb);
        stage_window(va, g0 + 4);
        tmem_ld_wait();
        stage_window(vb, g0 + 6);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(smem_u32(&s_tempty[buf]));          // this warp's share of the accumulator is drained
    }
  } else {
    // ===================== emitters: ring -> P8 =====================
    // warps 3, 7, 10, 11 (group 0) and 12, 13, 14, 15 (group 1); group g emits the windows with (running index) % 2 == g;
    // within a group warp -> 8-channel chunk, lane -> pixel (512-byte contiguous stores per plane).
    const int ei = (warp == 3) ? 0 : (warp == 7 ? 1 : warp - 8);     // 0..7
    const int grp = ei >> 2, ech = ei & 3, ep = lane;
    const float* S_ld = reinterpret_cast<const float*>(smem + w_bytes) + 8 * ech;
    const size_t oplane = (size_t)(a.AH + 2) * (a.AW + 2) * 8;
    const P8QGeom qg = P8QGeom::make(a.B, TCX_CO, a.AH, a.AW);
    float sc[8], sh[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) { sc[e] = s_scale[8 * ech + e]; sh[e] = s_shift[8 * ech + e]; }
    uint32_t gw = 0;
    for (int tile = blockIdx.x; tile < a.ntiles; tile += gridDim.x) {
      const int tx = tile % a.tiles_x, ty = (tile / a.tiles_x) % a.tiles_y, n = tile / (a.tiles_x * a.tiles_y);
      __nv_bfloat16* obase = a.out + ((size_t)n * (TCX_CO / 8) + ech) * 2 * oplane;
      __nv_bfloat16* qbase = a.outq ? a.outq + qg.slab(n, ech, 0) : nullptr;
      // (row, column) of this thread's output position in the tile, advanced by 32 per window without divisions
      int rr = ep >= 2 ? (ep - 2) / a.P : 0;                         // window 0: m = -2 + ep
      int cc = ep - 2 - rr * a.P;
      if (a.dbg & 1) { gw += TCX_N / 32; continue; }
#pragma unroll 1
      for (int w = 0; w < TCX_N / 32; ++w, ++gw) {
        if ((int)(gw & 1) == grp) {
          mbar_wait(smem_u32(&s_sfull[gw & 3]), (gw >> 2) & 1);
          // rows 32 gw - 2 and 32 gw - 1 hold columns of the previous window, staged by the OTHER stager set
          if (w > 0) mbar_wait(smem_u32(&s_sfull[(gw - 1) & 3]), ((gw - 1) >> 2) & 1);
          const int m = 32 * w - 2 + ep;
          const int ay = ty * a.R + rr, ax = tx * a.Wt + cc;
          if (m >= 0 && rr < a.R && cc < a.Wt && ay < a.AH && ax < a.AW && !(a.dbg & 16)) {
            const float* sp = S_ld + ((32 * gw - 2 + ep) & (TCX_S_ROWS - 1)) * TCX_S_PITCH;
            float f[8];
#pragma unroll
            for (int hv = 0; hv < 2; ++hv) {
              const float4 v0 = *reinterpret_cast<const float4*>(sp + 4 * hv);
              const float4 v1 = *reinterpret_cast<const float4*>(sp + 4 * hv + TCX_S_PLANE);
              const float4 v2 = *reinterpret_cast<const float4*>(sp + 4 * hv + 2 * TCX_S_PLANE);
              f[4 * hv + 0] = v0.x + v1.x + v2.x; f[4 * hv + 1] = v0.y + v1.y + v2.y;
              f[4 * hv + 2] = v0.z + v1.z + v2.z; f[4 * hv + 3] = v0.w + v1.w + v2.w;
            }
#pragma unroll
            for (int e = 0; e < 8; ++e) {
              f[e] = fmaf(f[e], sc[e], sh[e]);
              if (a.relu) f[e] = fmaxf(f[e], 0.f);
            }
            uint4 hi, lo;
            split8_fast(f, hi, lo);
            __nv_bfloat16* p = obase + ((size_t)(ay + 1) * (a.AW + 2) + (ax + 1)) * 8;
            if (!((a.dbg & 32) && hi.x != 0x12345678u)) {             // (timing experiment: no global stores)
              *reinterpret_cast<uint4*>(p) = hi;
              *reinterpret_cast<uint4*>(p + oplane) = lo;
              if (qbase) {
                __nv_bfloat16* pq = qbase + qg.pixel(ay, ax);
                *reinterpret_cast<uint4*>(pq) = hi;
                *reinterpret_cast<uint4*>(pq + qg.slab_elems()) = lo;
              }
            }
          }
          __syncwarp();
          if (lane == 0) mbar_arrive(smem_u32(&s_sempty[gw & 3]));
        }
        cc += 32;
        while (cc >= a.P) { cc -= a.P; ++rr; }
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

// ---- weight packing -> [kstage][dy][slot = hi j0 | hi j1 | lo j0 | lo j1][128][8] -------------------------
//   hi slots serve both K-halves of their MMA (LBO = 0) against [x_hi ; x_lo]; the two lo slots are the K-halves
//   of the third MMA against x_hi of both chunks.
__global__ void tcx_pack_weights_kernel(const float* __restrict__ w, __nv_bfloat16* __restrict__ wp, int CI, int nstage) {
  const int total = nstage * 3 * 4 * 128 * 8;
  for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += gridDim.x * blockDim.x) {
    int e = idx & 7;
    int r = idx >> 3;
    const int m = r & 127; r >>= 7;
    const int slot = r & 3; r >>= 2;
    const int dy = r % 3;
    const int ks = r / 3;
    const int g = m >> 5, co = m & 31;
    float v = 0.f;
    if (g < 3) {
      const int ci = ks * 16 + (slot & 1) * 8 + e;
      if (ci < CI) v = __ldg(w + ((size_t)co * CI + ci) * 9 + dy * 3 + g);
    }
    __nv_bfloat16 hi, lo;
    split_bf16(v, hi, lo);
    wp[idx] = (slot < 2) ? hi : lo;
  }
}

struct TcxPlan { int R, Wt, P, rows_alloc, rows_ld, tiles_x, tiles_y, stages; size_t smem; bool ok; };

static TcxPlan tcx_plan(int AH, int AW, int nk) {
  TcxPlan best = {}; best.ok = false;
  double best_cost = 1e300;
  const size_t fixed = (size_t)nk * TCX_WSTAGE_BYTES + (size_t)TCX_S_BYTES;
  for (int R = 1; R <= 32; ++R) {
    for (int P = 10; P <= 64; ++P) {                                   // 4P words <= 256 (tensor-map box limit)
      const int Wt = P - 2;
      if (R * P - 2 > TCX_OUT) break;
      if (Wt > AW + 7) break;
      int rows_alloc = (TCX_N + 2 * P + P - 1) / P;
      while ((rows_alloc * P * 16) % 128) ++rows_alloc;              // tensor copies want 128 B aligned planes
      const size_t stage = (size_t)4 * rows_alloc * P * 16;
      if (fixed + 2 * stage > (size_t)TC_SMEM_MAX) continue;          // (also keeps the unsigned difference below from wrapping)
      int stages = (int)((TC_SMEM_MAX - fixed) / stage);
      if (stages > TCX_MAX_STAGES) stages = TCX_MAX_STAGES;
      const int tx = (AW + Wt - 1) / Wt, ty = (AH + R - 1) / R;
      // every tile costs the same MMA time; loads (halo rows re-read) weigh in lightly
      const double cost = (double)tx * ty * (1.0 + 0.15 * (double)((R + 2) * P) / TCX_N);
      if (cost < best_cost) {
        best_cost = cost;
        best = {R, Wt, P, rows_alloc, R + 2, tx, ty, stages, fixed + (size_t)stages * stage, true};
      }
    }
  }
  return best;
}

static int launch_tcx(const TcxArgs& a, const TcxPlan& p, cudaStream_t st) {
  static bool attr_done[NCONV_MAX_DEVICES] = {};
  static int sm_count[NCONV_MAX_DEVICES] = {};
  int dev = 0, sms = 0;
  NCONV_RETURN_IF_ERR(cudaGetDevice(&dev));
  if (dev >= NCONV_MAX_DEVICES || !attr_done[dev]) {
    NCONV_RETURN_IF_ERR(cudaFuncSetAttribute(conv_tcx_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM_MAX));
    if (dev < NCONV_MAX_DEVICES) attr_done[dev] = true;
  }
  if (dev < NCONV_MAX_DEVICES) sms = sm_count[dev];
  if (sms == 0) {
    NCONV_RETURN_IF_ERR(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    if (dev < NCONV_MAX_DEVICES) sm_count[dev] = sms;
  }
  const int grid = a.ntiles < sms ? a.ntiles : sms;
  const int CI = 16 * (a.pairs[0] + a.pairs[1]);
  const double opx = (double)a.B * a.AH * a.AW;
  alignas(64) CUtensorMap tm[2];
  for (int si = 0; si < 2; ++si) {
    const int use = (a.src[si] && a.c8[si] > 0) ? si : 0;
    const int rc = tc_make_tmap(&tm[si], a.src[use], a.B * a.c8[use] * 2, a.AH + 2, a.AW + 2, a.P, a.rows_ld);
    if (rc) return rc;
  }
  char label[64];
  snprintf(label, sizeof(label), "tc_conv3x3 %d->%d @%dx%d", CI, TCX_CO, a.AH, a.AW);
  {
    LaunchScope ls(st, label, 2.0 * CI * TCX_CO * 9.0 * opx, 4.0 * (CI * opx + TCX_CO * opx * (a.outq ? 2.0 : 1.0)));
    conv_tcx_kernel<<<grid, TCX_THREADS, p.smem, st>>>(a, tm[0], tm[1]);
  }
  return (int)cudaGetLastError();
}

}  // namespace nconv
