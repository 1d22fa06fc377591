// P8 layout utilities (csrc/p8.cuh): fp32 NCHW <-> P8 conversion at the boundaries of the
// RGB-guided stack and in tests, and the zero ring of reused buffers.
#include "common.cuh"
#include "launch_count.h"
#include "nconv_b200.h"
#include "p8.cuh"

namespace nconv {

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

namespace nconv {
// zero row 0 and column 0 of every P8Q phase plane
__global__ void p8q_zero_ring_kernel(__nv_bfloat16* __restrict__ p, size_t planes, int Hq1, int Wq1) {
  const int per = Wq1 + Hq1 - 1;
  const size_t total = planes * per;
  const uint4 z = make_uint4(0, 0, 0, 0);
  for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
    int i = (int)(idx % per);
    size_t pl = idx / per;
    int y = (i < Wq1) ? 0 : i - Wq1 + 1, x = (i < Wq1) ? i : 0;
    *reinterpret_cast<uint4*>(p + (pl * Hq1 * Wq1 + (size_t)y * Wq1 + x) * 8) = z;
  }
}
}  // namespace nconv

extern "C" size_t nconv_b200_p8q_bytes(int B, int C, int H, int W) {
  if (B <= 0 || C <= 0 || H <= 0 || W <= 0 || ((H | W) & 1)) return 0;
  return P8QGeom::make(B, C, H, W).total_elems() * sizeof(__nv_bfloat16) + P8_TAIL_SLACK_BYTES;
}

extern "C" int nconv_b200_p8q_zero_ring(void* p8q, int B, int C, int H, int W, void* stream) {
  if (!p8q || B <= 0 || C <= 0 || H <= 0 || W <= 0 || ((H | W) & 1)) return NCONV_ERR_BAD_ARG;
  const P8QGeom g = P8QGeom::make(B, C, H, W);
  const size_t planes = (size_t)B * g.C8 * 2 * 4;
  const size_t total = planes * (g.Wq1 + g.Hq1 - 1);
  int blocks = (int)((total + 255) / 256); if (blocks > 148 * 8) blocks = 148 * 8;
  {
    LaunchScope ls((cudaStream_t)stream, "p8q_zero_ring", 0.0, 16.0 * total);
    p8q_zero_ring_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>((__nv_bfloat16*)p8q, planes, g.Hq1, g.Wq1);
  }
  return (int)cudaGetLastError();
}

namespace nconv {
__global__ void p8q_pack_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ p, int B, int C, int H, int W) {
  const P8QGeom g = P8QGeom::make(B, C, H, W);
  const size_t total = (size_t)B * g.C8 * 4 * g.Hq1 * g.Wq1;
  for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
    int xq = (int)(idx % g.Wq1);
    size_t r = idx / g.Wq1;
    int yq = (int)(r % g.Hq1); r /= g.Hq1;
    int ph = (int)(r & 3); r >>= 2;
    int ch = (int)(r % g.C8);
    int n = (int)(r / g.C8);
    const int yy = 2 * (yq - 1) + (ph >> 1), xx = 2 * (xq - 1) + (ph & 1);
    const bool inside = yq >= 1 && xq >= 1;
    float f[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int c = ch * 8 + e;
      f[e] = (inside && c < C) ? __ldg(x + (((size_t)n * C + c) * H + yy) * W + xx) : 0.f;
    }
    uint4 hi, lo;
    split8(f, hi, lo);
    const size_t off = (size_t)ph * g.phase_elems() + ((size_t)yq * g.Wq1 + xq) * 8;
    *reinterpret_cast<uint4*>(p + g.slab(n, ch, 0) + off) = hi;
    *reinterpret_cast<uint4*>(p + g.slab(n, ch, 1) + off) = lo;
  }
}
}  // namespace nconv

extern "C" int nconv_b200_p8q_pack(const float* x, void* p8q, int B, int C, int H, int W, void* stream) {
  if (!x || !p8q || B <= 0 || C <= 0 || H <= 0 || W <= 0 || ((H | W) & 1)) return NCONV_ERR_BAD_ARG;
  const P8QGeom g = P8QGeom::make(B, C, H, W);
  const size_t total = (size_t)B * g.C8 * 4 * g.Hq1 * g.Wq1;
  int blocks = (int)((total + 255) / 256); if (blocks > 148 * 16) blocks = 148 * 16;
  {
    LaunchScope ls((cudaStream_t)stream, "p8q_pack", 0.0, 8.0 * B * C * H * W);
    p8q_pack_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(x, (__nv_bfloat16*)p8q, B, C, H, W);
  }
  return (int)cudaGetLastError();
}
