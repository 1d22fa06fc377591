// "P8" activation layout of the RGB-guided stack (SETP2, /root/reference/models/step2.py) on B200.
//
// A [B,C,H,W] fp32 activation is held as split bf16 (hi = bf16(v), lo = bf16(v - hi): 16 mantissa
// bits, the precision the tcgen05 bf16x3 contraction consumes) in pixel-major 8-channel chunks with
// a one-pixel ring of zeros:
//
//     p8[n][chunk = c/8][part = hi|lo][y = 0..H+1][x = 0..W+1][c%8]      (bf16, 16 B per pixel)
//
// * 16 B per (pixel, chunk, part) is one row of a tcgen05 K-major no-swizzle core matrix, so a run
//   of consecutive pixels IS a canonical A-operand tile and "shift by one pixel" is "+16 B on the
//   descriptor start address": the 3x3 taps become nine descriptor offsets into one halo tile
//   (implicit im2col without materialising patches).
// * the zero ring is the convolution's zero padding, so tile loads never test bounds; kernels write
//   interior pixels only and the ring is written once per forward by p8_zero_ring.
// * channels beyond C in the last chunk are zero.
// * same bytes per activation as fp32 (2 x bf16), so HBM traffic is unchanged.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace nconv {

struct P8Geom {
  int B, C8, H, W, Hp, Wp;
  __host__ __device__ static P8Geom make(int B, int C, int H, int W) {
    P8Geom g; g.B = B; g.C8 = (C + 7) / 8; g.H = H; g.W = W; g.Hp = H + 2; g.Wp = W + 2; return g;
  }
  // element (bf16) offsets
  __host__ __device__ size_t plane_elems() const { return (size_t)Hp * Wp * 8; }
  __host__ __device__ size_t plane(int n, int chunk, int part) const {
    return (((size_t)n * C8 + chunk) * 2 + part) * plane_elems();
  }
  __host__ __device__ size_t pixel(int y, int x) const { return ((size_t)(y + 1) * Wp + (x + 1)) * 8; }  // interior coords
  __host__ __device__ size_t total_elems() const { return (size_t)B * C8 * 2 * plane_elems(); }
};


// "P8Q": the same activation split into its four spatial phases, for stride-2 consumers
// (RGBEncoder's 3x3/s2 and 1x1/s2 convolutions, step2.py:134-154).  H and W even:
//
//     p8q[n][chunk][part][phase = (y&1)*2 + (x&1)][1 + y/2][1 + x/2][c%8],  planes (H/2+1) x (W/2+1)
//
// Row 0 and column 0 of every plane are zero (the pad-1 halo of a stride-2 3x3 only ever reaches
// one pixel up/left in phase space), so a stride-2 tap is again "phase plane + flat pixel shift".
struct P8QGeom {
  int B, C8, Hq1, Wq1;
  __host__ __device__ static P8QGeom make(int B, int C, int H, int W) {
    P8QGeom g; g.B = B; g.C8 = (C + 7) / 8; g.Hq1 = H / 2 + 1; g.Wq1 = W / 2 + 1; return g;
  }
  __host__ __device__ size_t phase_elems() const { return (size_t)Hq1 * Wq1 * 8; }
  __host__ __device__ size_t slab_elems() const { return 4 * phase_elems(); }
  __host__ __device__ size_t slab(int n, int chunk, int part) const { return (((size_t)n * C8 + chunk) * 2 + part) * slab_elems(); }
  __host__ __device__ size_t pixel(int y, int x) const {   // full-resolution interior coords
    return (size_t)((y & 1) * 2 + (x & 1)) * phase_elems() + ((size_t)((y >> 1) + 1) * Wq1 + ((x >> 1) + 1)) * 8;
  }
  __host__ __device__ size_t total_elems() const { return (size_t)B * C8 * 2 * slab_elems(); }
};

constexpr size_t P8_TAIL_SLACK_BYTES = 8192;   // tile loads may over-read one staged row past the end

__device__ __forceinline__ void split_bf16(float v, __nv_bfloat16& hi, __nv_bfloat16& lo) {
  hi = __float2bfloat16_rn(v);
  lo = __float2bfloat16_rn(v - __bfloat162float(hi));
}

// 8 floats -> (hi, lo) 16-byte vectors
__device__ __forceinline__ void split8(const float* v, uint4& hi, uint4& lo) {
  __nv_bfloat16 h[8], l[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) split_bf16(v[i], h[i], l[i]);
  hi = *reinterpret_cast<uint4*>(h);
  lo = *reinterpret_cast<uint4*>(l);
}

// Same result as split8 with packed conversions (cvt.rn.bf16x2.f32: two values per instruction; the bf16 -> fp32
// back-conversion is a shift / mask): ~3 instructions per value instead of ~5.
__device__ __forceinline__ void split8_fast(const float* v, uint4& hi, uint4& lo) {
  uint32_t h[4], l[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const __nv_bfloat162 hp = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
    const uint32_t hb = *reinterpret_cast<const uint32_t*>(&hp);
    const float r0 = v[2 * i] - __uint_as_float(hb << 16);
    const float r1 = v[2 * i + 1] - __uint_as_float(hb & 0xffff0000u);
    const __nv_bfloat162 lp = __floats2bfloat162_rn(r0, r1);
    h[i] = hb;
    l[i] = *reinterpret_cast<const uint32_t*>(&lp);
  }
  hi = make_uint4(h[0], h[1], h[2], h[3]);
  lo = make_uint4(l[0], l[1], l[2], l[3]);
}

__device__ __forceinline__ void split4_fast(const float* v, uint2& hi, uint2& lo) {
  uint32_t h[2], l[2];
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const __nv_bfloat162 hp = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
    const uint32_t hb = *reinterpret_cast<const uint32_t*>(&hp);
    const __nv_bfloat162 lp = __floats2bfloat162_rn(v[2 * i] - __uint_as_float(hb << 16), v[2 * i + 1] - __uint_as_float(hb & 0xffff0000u));
    h[i] = hb;
    l[i] = *reinterpret_cast<const uint32_t*>(&lp);
  }
  hi = make_uint2(h[0], h[1]);
  lo = make_uint2(l[0], l[1]);
}

__device__ __forceinline__ void join8(const uint4& hi, const uint4& lo, float* v) {
  const __nv_bfloat16* h = reinterpret_cast<const __nv_bfloat16*>(&hi);
  const __nv_bfloat16* l = reinterpret_cast<const __nv_bfloat16*>(&lo);
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] = __bfloat162float(h[i]) + __bfloat162float(l[i]);
}

}  // namespace nconv
