// "N8" activation layout of the tensor-core NConv layers (DNET, /root/reference/models/step1.py:38-94).
//
// An 8-channel (data, confidence) pair is stored as the two streams NConv2d correlates with its weights
// (step1.py:119-121: conf and data*conf), each split to bf16 hi/lo like P8 (p8.cuh), pixel-major with a 2-pixel ring
// of zeros (the zero padding of a 5x5 / pad 2 layer; a 3x3 / pad 1 layer uses the inner ring pixel only):
//
//     n8[n][stream = xc | c][part = hi | lo][y = 0..H+3][x = 0..W+3][8 channels]      bf16, 16 B per (pixel, plane)
//
// 16 B per pixel per plane is one row of a tcgen05 K-major no-swizzle core matrix, so 128 consecutive pixels of a
// staged tile are one A operand and a vertical tap is a start-address shift of one staged row.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace nconv {

constexpr int N8_RING = 2;

struct N8Geom {
  int B, H, W, Hp, Wp;
  __host__ __device__ static N8Geom make(int B, int H, int W) {
    N8Geom g; g.B = B; g.H = H; g.W = W; g.Hp = H + 2 * N8_RING; g.Wp = W + 2 * N8_RING; return g;
  }
  __host__ __device__ size_t plane_elems() const { return (size_t)Hp * Wp * 8; }
  // plane index: (n*2 + stream)*2 + part
  __host__ __device__ size_t plane(int n, int stream, int part) const { return (((size_t)n * 2 + stream) * 2 + part) * plane_elems(); }
  __host__ __device__ size_t pixel(int y, int x) const { return ((size_t)(y + N8_RING) * Wp + (x + N8_RING)) * 8; }   // interior coords
  __host__ __device__ size_t total_elems() const { return (size_t)B * 4 * plane_elems(); }
};

constexpr size_t N8_TAIL_SLACK_BYTES = 1024;

}  // namespace nconv
