// GPU input pipeline: the per-sample host work of the reference's loaders, moved behind the H2D copy
// so that a batch crosses PCIe as the bytes the files hold (uint8 / uint16) instead of fp32:
//   * RGB: cv2.imread -> torch.FloatTensor(...).permute(2, 0, 1)      dataset/nyuloader.py:69, kittiloader.py:63
//          uint8 [B,H,W,3] (BGR, HWC)  ->  fp32 [B,3,H,W]
//   * depth / gt: cv2.imread(IMREAD_GRAYSCALE).astype(float32) / 256  kittiloader.py:66-76   (uint8 or uint16 PNG, or fp32 .npy)
//          [B,H,W] -> fp32 [B,1,H,W] * scale
//   * sparsifier: depth + depth*noise on the noisy points, then the mask multiply (or the zeroed random points, which is
//          a 0/1 mask as well)                                        nyuloader.py:83-124
//          out = raw * (1 + noise) * mask         noise: fp32 or NULL, mask: uint8 0/1 or NULL
// All three are elementwise and HBM-bound; which pixels get noise / get dropped is the caller's random draw (the reference
// uses torch.randperm on the host), passed in as the noise and mask tensors.
#include "common.cuh"
#include "launch_count.h"
#include "nconv_b200.h"

namespace nconv {

__global__ void rgb_u8_hwc_to_f32_chw_kernel(const uint8_t* __restrict__ in, float* __restrict__ out, int B, int H, int W) {
  const long long px = (long long)B * H * W;
  const long long hw = (long long)H * W;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < px; i += (long long)gridDim.x * blockDim.x) {
    const long long n = i / hw, p = i - n * hw;
    const uint8_t* s = in + i * 3;
    float* d = out + n * 3 * hw + p;
    d[0] = (float)s[0]; d[hw] = (float)s[1]; d[2 * hw] = (float)s[2];
  }
}

template <typename T>
__global__ void depth_decode_kernel(const T* __restrict__ in, float* __restrict__ out, long long n, float scale) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    out[i] = (float)in[i] * scale;
}

__global__ void sparsify_kernel(const float* __restrict__ raw, const float* __restrict__ noise, const uint8_t* __restrict__ mask,
                                float* __restrict__ out, long long n) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    float d = raw[i];
    if (noise) d = d + d * noise[i];                    // nyuloader.py:107 `depth[idx] += depth[idx] * noise`
    if (mask) d = d * (float)mask[i];                   // :113 `depth * mask`  /  :121-123 zeroed points
    out[i] = d;
  }
}

static int grid_for(long long n) {
  long long b = (n + 255) / 256;
  return (int)(b > 148 * 16 ? 148 * 16 : (b < 1 ? 1 : b));
}

}  // namespace nconv

using namespace nconv;

extern "C" int nconv_b200_rgb_u8_to_f32(const uint8_t* hwc, float* chw, int B, int H, int W, void* stream) {
  if (!hwc || !chw || B <= 0 || H <= 0 || W <= 0) return NCONV_ERR_BAD_ARG;
  const long long px = (long long)B * H * W;
  {
    LaunchScope ls((cudaStream_t)stream, "rgb_u8_to_f32", 0.0, 15.0 * px);
    rgb_u8_hwc_to_f32_chw_kernel<<<grid_for(px), 256, 0, (cudaStream_t)stream>>>(hwc, chw, B, H, W);
  }
  return (int)cudaGetLastError();
}

extern "C" int nconv_b200_depth_decode(const void* raw, int elem_bytes, float* out, long long n, float scale, void* stream) {
  if (!raw || !out || n <= 0) return NCONV_ERR_BAD_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  LaunchScope ls(st, "depth_decode", 0.0, (4.0 + elem_bytes) * n);
  if (elem_bytes == 1) depth_decode_kernel<uint8_t><<<grid_for(n), 256, 0, st>>>((const uint8_t*)raw, out, n, scale);
  else if (elem_bytes == 2) depth_decode_kernel<uint16_t><<<grid_for(n), 256, 0, st>>>((const uint16_t*)raw, out, n, scale);
  else if (elem_bytes == 4) depth_decode_kernel<float><<<grid_for(n), 256, 0, st>>>((const float*)raw, out, n, scale);
  else return NCONV_ERR_UNSUPPORTED;
  return (int)cudaGetLastError();
}

extern "C" int nconv_b200_sparsify(const float* raw, const float* noise, const uint8_t* mask, float* out, long long n, void* stream) {
  if (!raw || !out || n <= 0) return NCONV_ERR_BAD_ARG;
  {
    LaunchScope ls((cudaStream_t)stream, "sparsify", 0.0, (8.0 + (noise ? 4.0 : 0.0) + (mask ? 1.0 : 0.0)) * n);
    sparsify_kernel<<<grid_for(n), 256, 0, (cudaStream_t)stream>>>(raw, noise, mask, out, n);
  }
  return (int)cudaGetLastError();
}
