// Weight gradient of the dense 3x3 / stride-1 convolutions of the RGB-guided stack on tcgen05
// (/root/reference/train_step2.py:60-66 differentiates models/step2.py:178-195, 216-236, 290-297 with autograd; this is
// the dL/dW half of a conv layer's backward):
//
//     gW[co][ci][u][v] = sum over (n, y, x) of  g[n][co][y][x] * x[n][ci][y+u-1][x+v-1]
//
// As a GEMM the contraction runs over PIXELS, so both operands are "MN-major" for the tensor core: in the P8 layout
// (p8.cuh) a pixel's 8 channels are 16 contiguous bytes and consecutive pixels follow at 16 B, i.e. 8 pixels x 8 channels
// is one 128-byte core matrix with the pixel (K) index on its rows.  tools/micro/mn_major_probe.cu established the
// descriptor convention on sm_100a: a_major = b_major = MN in the instruction descriptor, LBO = stride between the two
// 8-pixel core matrices of a K = 16 step (128 B), SBO = stride between 8-channel groups (= the P8 plane stride).
//
//   D[128 rows = (ci chunk, hi|lo) x 8][N = Co]  +=  A[x planes, 16 pixels at tap shift] * B[g planes, 16 pixels]^T
//
// * A's 128 rows are 64 input channels x {hi, lo}: P8 orders planes as (chunk, part), so the plane stride is uniform and
//   one MMA multiplies both halves of the split; B is g_hi, a second MMA takes g_lo: (x_hi + x_lo) * (g_hi + g_lo),
//   all four partial products in fp32.  The epilogue adds row pairs (hi, lo) with one shuffle.
// * a tap (u, v) is a start-address shift of the staged x tile (pitch 34 pixels); g tiles are 32 pixels wide and staged
//   densely, so a 16-pixel K step never crosses a row in either operand and no halo column of g needs masking.
// * the nine taps need 9 x Co accumulator columns (576 at Co = 64 > 512 of TMEM): taps are split in two groups (5 + 4),
//   input channels in blocks of 64, each (tap group, channel block) handled by its own set of persistent CTAs
//   (blockIdx.y), which accumulate over ALL their pixel tiles in TMEM and drain once at the end into per-CTA partial
//   sums; a second kernel adds the partials in a fixed order (deterministic) and writes gW in [Co][Ci][3][3] order.
#include "common.cuh"
#include "launch_count.h"
#include "nconv_b200.h"
#include "p8.cuh"
#include "tc_ptx.cuh"
#include <cuda.h>
#include <cstdio>
#include <cstdlib>

namespace nconv {

constexpr int WG_THREADS = 192;             // warp 0: producer, warp 1: MMA issuer, warps 2-5: final drain
constexpr int WG_R = 4, WG_WT = 32;         // output pixels per tile: 4 rows x 32 columns = 8 K steps
constexpr int WG_XP = WG_WT + 2;            // staged x pitch (pixels)
constexpr int WG_XROWS = WG_R + 2;
constexpr int WG_XPL = WG_XROWS * WG_XP * 16;      // bytes of one staged x plane (3264)
constexpr int WG_GPL = WG_R * WG_WT * 16;          // bytes of one staged g plane (2048)
constexpr int WG_XBLK = 16 * WG_XPL;               // 64 channels x {hi, lo}
constexpr int WG_STAGES = 2;

struct WgArgs {
  float* partial;               // [grid.x][ngroup(blockIdx.y)][taps_in_group <= 5][64 ci][CO]
  int B, CI, CO, H, W;
  int tiles_x, tiles_y, ntiles;
  int ci_blocks;                // CI / 64 (at least 1; CI = 32 uses one block with 8 live planes)
  int x_planes;                 // live x planes per block: min(CI, 64) / 8 * 2
};

template <int CO>
__global__ void __launch_bounds__(WG_THREADS, 1) conv_wgrad_tc_kernel(const __grid_constant__ WgArgs a, const __grid_constant__ CUtensorMap tmx,
                                                                      const __grid_constant__ CUtensorMap tmg) {
  constexpr int GBLK = CO / 8 * 2 * WG_GPL;            // staged g tile: all Co chunks x {hi, lo}
  constexpr int STAGE = WG_XBLK + GBLK;
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t s_full[WG_STAGES], s_empty[WG_STAGES], s_done;
  __shared__ uint32_t s_tmem_base;
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
  const uint32_t smem_base = smem_u32(smem);
  const int tg = blockIdx.y & 1, cb = blockIdx.y >> 1;     // tap group (0: taps 0-4, 1: taps 5-8), input-channel block
  const int tap0 = tg * 5, ntap = tg ? 4 : 5;

  if (threadIdx.x == 0) {
    for (int s = 0; s < WG_STAGES; ++s) { mbar_init(smem_u32(&s_full[s]), 1); mbar_init(smem_u32(&s_empty[s]), 1); }
    mbar_init(smem_u32(&s_done), 1);
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
    // ===================== producer: two tensor copies per tile =====================
    uint32_t stage = 0, phase = 0;
    for (int tile = blockIdx.x; tile < a.ntiles; tile += gridDim.x) {
      const int tx = tile % a.tiles_x, ty = (tile / a.tiles_x) % a.tiles_y, n = tile / (a.tiles_x * a.tiles_y);
      const int y0 = ty * WG_R, x0 = tx * WG_WT;
      mbar_wait_warp(smem_u32(&s_empty[stage]), phase ^ 1);
      const uint32_t full = smem_u32(&s_full[stage]);
      const uint32_t sa = smem_base + stage * STAGE;
      if (lane == 0) {
        mbar_arrive_expect_tx(full, (uint32_t)a.x_planes * WG_XPL + GBLK);
        // x: planes (n*C8 + 8*cb + chunk)*2 + part, rows y0-1 .. y0+R (padded coords y0 .. y0+R+1), columns x0-1 .. x0+Wt
        asm volatile(
            "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
            ::"r"(sa), "l"(reinterpret_cast<uint64_t>(&tmx)), "r"(4 * x0), "r"(y0), "r"((n * (a.CI / 8) + 8 * cb) * 2), "r"(full)
            : "memory");
        // g: all Co planes of image n, rows y0 .. y0+R-1 (padded +1), columns x0 .. x0+Wt-1 (padded +1)
        asm volatile(
            "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
            ::"r"(sa + WG_XBLK), "l"(reinterpret_cast<uint64_t>(&tmg)), "r"(4 * (x0 + 1)), "r"(y0 + 1), "r"(n * (CO / 8) * 2), "r"(full)
            : "memory");
      }
      __syncwarp();
      if (++stage == WG_STAGES) { stage = 0; phase ^= 1; }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (warp-convergent) =====================
    // idesc: c = f32, a = b = bf16, a_major = b_major = MN (bits 15, 16), N = CO, M = 128
    constexpr uint32_t IDESC = (1u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | (1u << 16) | ((uint32_t)(CO >> 3) << 17) | ((128u >> 4) << 24);
    constexpr uint32_t LBO16 = 128u >> 4;                       // K direction: next 8 pixels
    constexpr uint64_t A_HI = ((uint64_t)((WG_XPL >> 4) | (1u << 14))) << 32;          // SBO = x plane stride
    constexpr uint64_t B_HI = ((uint64_t)(((2 * WG_GPL) >> 4) | (1u << 14))) << 32;    // SBO = 2 g planes (same part of the next chunk)
    uint32_t stage = 0, phase = 0, first = 1;
    for (int tile = blockIdx.x; tile < a.ntiles; tile += gridDim.x) {
      mbar_wait_warp(smem_u32(&s_full[stage]), phase);
      tc_fence_after();
      const uint32_t xs = (smem_base + stage * STAGE) >> 4;
      const uint32_t gs = xs + (WG_XBLK >> 4);
#pragma unroll 1
      for (int ks = 0; ks < WG_R * WG_WT / 16; ++ks) {
        const uint32_t r = ks >> 1, c0 = (ks & 1) * 16;
        const uint32_t g_hi = gs + (r * WG_WT + c0), g_lo = g_hi + (WG_GPL >> 4);
#pragma unroll
        for (int t = 0; t < 5; ++t) {
          if (t < ntap) {
            const int tap = tap0 + t, u = tap / 3, v = tap % 3;
            const uint32_t ax = xs + ((r + u) * WG_XP + c0 + v);
            const uint64_t ad = A_HI | (uint64_t)(ax | (LBO16 << 16));
            const uint32_t d = tmem_base + t * CO;
            umma_bf16_ws(d, ad, B_HI | (uint64_t)(g_hi | (LBO16 << 16)), IDESC, first ? 0u : 1u);
            umma_bf16_ws(d, ad, B_HI | (uint64_t)(g_lo | (LBO16 << 16)), IDESC, 1u);
          }
        }
        first = 0;
      }
      umma_commit_ws(smem_u32(&s_empty[stage]));
      if (++stage == WG_STAGES) { stage = 0; phase ^= 1; }
    }
    umma_commit_ws(smem_u32(&s_done));
  } else {
    // ===================== final drain (warps 2..5): TMEM -> this CTA's partial sums =====================
    const int q = warp & 3;
    mbar_wait(smem_u32(&s_done), 0);
    tc_fence_after();
    // row = 32 q + lane: plane p = row / 8 = (chunk j, part), channel within the chunk = lane % 8; hi rows have even p
    const int p = 4 * q + (lane >> 3), c = lane & 7;
    const int ci_local = 8 * (p >> 1) + c;                       // 0..63 within the block
    const bool hi_row = (p & 1) == 0;
    float* out = a.partial + (((size_t)blockIdx.x * gridDim.y + blockIdx.y) * 5) * 64 * CO;
    for (int t = 0; t < ntap; ++t) {
#pragma unroll
      for (int cb32 = 0; cb32 < CO / 32; ++cb32) {
        uint32_t v[32];
        tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + t * CO + cb32 * 32, v);
        tmem_ld_wait();
        float f[32];
#pragma unroll
        for (int e = 0; e < 32; ++e) {
          const float x = __uint_as_float(v[e]);
          f[e] = x + __shfl_down_sync(0xffffffffu, x, 8);        // hi row + the lo row 8 lanes up
        }
        if (hi_row) {
          float* o = out + ((size_t)t * 64 + ci_local) * CO + cb32 * 32;
#pragma unroll
          for (int e = 0; e < 32; e += 4) *reinterpret_cast<float4*>(o + e) = make_float4(f[e], f[e + 1], f[e + 2], f[e + 3]);
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

// gw[co][ci][tap] = sum over CTAs of partial[cta][group = (cb, tg)][t][ci_local][co], fixed order (deterministic)
__global__ void conv_wgrad_tc_reduce_kernel(const float* __restrict__ partial, float* __restrict__ gw, int nx, int ny, int CI, int CO) {
  const int total = CO * CI * 9;
  for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += gridDim.x * blockDim.x) {
    // idx enumerates (tap, ci, co) with co fastest so that the partial reads are coalesced
    const int co = idx % CO, ci = (idx / CO) % CI, tap = idx / (CO * CI);
    const int tg = tap >= 5 ? 1 : 0, t = tap - 5 * tg, cb = ci >> 6, cl = ci & 63;
    const int y = cb * 2 + tg;
    double s = 0.0;
    for (int x = 0; x < nx; ++x) s += (double)partial[((((size_t)x * ny + y) * 5 + t) * 64 + cl) * CO + co];
    gw[((size_t)co * CI + ci) * 9 + tap] = (float)s;
  }
}

typedef CUresult (*WgEncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                               const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                               CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// P8 tensor seen as [planes][rows][pitch * 4 words]; box [box_px * 4 words][box_rows][box_planes]
static int wg_make_tmap(CUtensorMap* tm, const void* base, int planes, int rows, int pitch, int box_px, int box_rows, int box_planes) {
  static WgEncodeFn enc = nullptr;
  if (!enc) {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qr;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qr) != cudaSuccess || !fn) return NCONV_ERR_UNSUPPORTED;
    enc = (WgEncodeFn)fn;
  }
  const cuuint64_t gdim[3] = {(cuuint64_t)pitch * 4, (cuuint64_t)rows, (cuuint64_t)planes};
  const cuuint64_t gstr[2] = {(cuuint64_t)pitch * 16, (cuuint64_t)rows * pitch * 16};
  const cuuint32_t box[3] = {(cuuint32_t)box_px * 4, (cuuint32_t)box_rows, (cuuint32_t)box_planes};
  const cuuint32_t estr[3] = {1, 1, 1};
  const CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_UINT32, 3, const_cast<void*>(base), gdim, gstr, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : NCONV_ERR_BAD_ARG;
}

struct WgPlan { int nx, ny, tiles_x, tiles_y, ntiles, ci_blocks, x_planes; };

static bool wg_plan(int B, int CI, int CO, int H, int W, WgPlan* p) {
  if (!(CO == 32 || CO == 64) || !(CI == 32 || (CI % 64 == 0 && CI <= 256)) || B <= 0 || H <= 0 || W <= 0) return false;
  p->ci_blocks = CI <= 64 ? 1 : CI / 64;
  p->x_planes = (CI < 64 ? CI : 64) / 8 * 2;
  p->ny = 2 * p->ci_blocks;
  p->tiles_x = (W + WG_WT - 1) / WG_WT; p->tiles_y = (H + WG_R - 1) / WG_R;
  p->ntiles = p->tiles_x * p->tiles_y * B;
  int nx = 148 / p->ny;
  if (nx < 1) nx = 1;
  if (nx > p->ntiles) nx = p->ntiles;
  p->nx = nx;
  return true;
}

}  // namespace nconv

using namespace nconv;

extern "C" size_t nconv_b200_conv_wgrad_tc_ws_bytes(int B, int CI, int CO, int H, int W) {
  WgPlan p;
  if (!wg_plan(B, CI, CO, H, W, &p)) return 0;
  return (size_t)p.nx * p.ny * 5 * 64 * CO * sizeof(float);
}

extern "C" int nconv_b200_conv_wgrad_tc(const void* x_p8, const void* g_p8, float* gw, void* workspace, size_t workspace_bytes, int B,
                                        int CI, int CO, int H, int W, void* stream) {
  if (!x_p8 || !g_p8 || !gw || !workspace) return NCONV_ERR_BAD_ARG;
  WgPlan p;
  if (!wg_plan(B, CI, CO, H, W, &p)) return NCONV_ERR_UNSUPPORTED;
  if (workspace_bytes < nconv_b200_conv_wgrad_tc_ws_bytes(B, CI, CO, H, W)) return NCONV_ERR_WORKSPACE;
  cudaStream_t st = (cudaStream_t)stream;
  WgArgs a = {};
  a.partial = (float*)workspace;
  a.B = B; a.CI = CI; a.CO = CO; a.H = H; a.W = W;
  a.tiles_x = p.tiles_x; a.tiles_y = p.tiles_y; a.ntiles = p.ntiles; a.ci_blocks = p.ci_blocks; a.x_planes = p.x_planes;
  alignas(64) CUtensorMap tmx, tmg;
  int rc = wg_make_tmap(&tmx, x_p8, B * (CI / 8) * 2, H + 2, W + 2, WG_XP, WG_XROWS, p.x_planes);
  if (rc) return rc;
  rc = wg_make_tmap(&tmg, g_p8, B * (CO / 8) * 2, H + 2, W + 2, WG_WT, WG_R, CO / 8 * 2);
  if (rc) return rc;
  const size_t smem = (size_t)WG_STAGES * (WG_XBLK + CO / 8 * 2 * WG_GPL);
  static bool attr_done[NCONV_MAX_DEVICES][2] = {};
  int dev = 0;
  NCONV_RETURN_IF_ERR(cudaGetDevice(&dev));
  const int ki = CO == 64 ? 1 : 0;
  if (dev >= NCONV_MAX_DEVICES || !attr_done[dev][ki]) {
    if (CO == 64) NCONV_RETURN_IF_ERR(cudaFuncSetAttribute(conv_wgrad_tc_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    else NCONV_RETURN_IF_ERR(cudaFuncSetAttribute(conv_wgrad_tc_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    if (dev < NCONV_MAX_DEVICES) attr_done[dev][ki] = true;
  }
  const double px = (double)B * H * W;
  char label[64];
  snprintf(label, sizeof(label), "tc_conv_wgrad k3 s1 %d->%d @%dx%d", CI, CO, H, W);
  {
    LaunchScope ls(st, label, 2.0 * CI * CO * 9.0 * px, 4.0 * (CI + CO) * px);
    if (CO == 64) conv_wgrad_tc_kernel<64><<<dim3(p.nx, p.ny), WG_THREADS, smem, st>>>(a, tmx, tmg);
    else conv_wgrad_tc_kernel<32><<<dim3(p.nx, p.ny), WG_THREADS, smem, st>>>(a, tmx, tmg);
  }
  NCONV_RETURN_IF_ERR(cudaGetLastError());
  {
    LaunchScope ls(st, "tc_conv_wgrad_reduce", 0.0, 0.0);
    const int total = CO * CI * 9;
    conv_wgrad_tc_reduce_kernel<<<(total + 255) / 256, 256, 0, st>>>(a.partial, gw, p.nx, p.ny, CI, CO);
  }
  return (int)cudaGetLastError();
}
