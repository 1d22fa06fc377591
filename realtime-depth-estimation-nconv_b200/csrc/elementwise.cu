// Small stand-alone stages of the NConv path: EnforcePos, threshold, 2x2 max-pool fwd/bwd,
// nearest-upsample+concat fwd/bwd, align-corners bilinear resize, border zeroing, BN folding.
// In inference most of these are fused into the NConv kernels (nconv_fwd.cu); the stand-alone
// versions serve the training graph and the per-stage parity tests.
#include "common.cuh"
#include "nconv_b200.h"
#include "launch_count.h"
#include <cstdio>
#include <mutex>
#include <vector>

namespace nconv {

std::atomic<unsigned long long> g_launch_count{0};
std::atomic<int> g_profile_on{0};

// ---- optional per-launch event timing (bench.py roofline leg) ------------------------------
struct ProfRecord {
  char label[64];
  double flops, bytes;
  cudaEvent_t a, b;
};
static std::mutex g_prof_mu;
static std::vector<ProfRecord*> g_prof;

void profile_open(cudaStream_t st, const char* label, double flops, double bytes, void** token) {
  ProfRecord* r = new ProfRecord;
  snprintf(r->label, sizeof(r->label), "%s", label);
  r->flops = flops; r->bytes = bytes;
  if (cudaEventCreate(&r->a) != cudaSuccess || cudaEventCreate(&r->b) != cudaSuccess) { delete r; return; }
  cudaEventRecord(r->a, st);
  *token = r;
}
void profile_close(cudaStream_t st, void* token) {
  ProfRecord* r = (ProfRecord*)token;
  cudaEventRecord(r->b, st);
  std::lock_guard<std::mutex> lk(g_prof_mu);
  g_prof.push_back(r);
}
static void profile_clear() {
  std::lock_guard<std::mutex> lk(g_prof_mu);
  for (ProfRecord* r : g_prof) { cudaEventDestroy(r->a); cudaEventDestroy(r->b); delete r; }
  g_prof.clear();
}

static inline int grid_for(long long n, int block = 256) {
  long long g = (n + block - 1) / block;
  long long cap = 148LL * 16;
  return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

// ---- EnforcePos (step1.py:190-212) -----------------------------------------------------
__global__ void enforce_pos_kernel(float* __restrict__ w, long long n, int mode, int group) {
  long long i0 = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  long long stride = (long long)gridDim.x * blockDim.x;
  if (mode == 3) {  // softmax over each contiguous group of `group` elements (step1.py:199-203)
    long long ng = n / group;
    for (long long g = i0; g < ng; g += stride) {
      float* p = w + g * group;
      float m = -INFINITY;
      for (int j = 0; j < group; ++j) m = fmaxf(m, p[j]);
      float s = 0.f;
      for (int j = 0; j < group; ++j) s += expf(p[j] - m);
      for (int j = 0; j < group; ++j) p[j] = expf(p[j] - m) / s;
    }
    return;
  }
  for (long long i = i0; i < n; i += stride) {
    float v = w[i];
    if (mode == 0) v = softplus_beta10(v);
    else if (mode == 1) v = expf(v);
    else v = 1.f / (1.f + expf(-v));
    w[i] = v;
  }
}

// ---- threshold (step1.py:53) -------------------------------------------------------------
__global__ void confidence_mask_kernel(const float* __restrict__ s, float* __restrict__ c,
                                       long long n, float thr) {
  long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += stride)
    c[i] = (__ldg(s + i) > thr) ? 1.f : 0.f;
}

// ---- 2x2/2 max-pool (step1.py:62-74) -----------------------------------------------------
__global__ void maxpool2x2_fwd_kernel(const float* __restrict__ x, float* __restrict__ y,
                                      int planes, int H, int W, int HP, int WP) {
  long long total = (long long)planes * HP * WP;
  long long stride = (long long)gridDim.x * blockDim.x;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total; idx += stride) {
    int px = (int)(idx % WP);
    long long r = idx / WP;
    int py = (int)(r % HP);
    long long pl = r / HP;
    const float* p = x + (pl * H + 2 * py) * W + 2 * px;
    float m = -INFINITY;
    m = pool_max(m, __ldg(p)); m = pool_max(m, __ldg(p + 1));
    m = pool_max(m, __ldg(p + W)); m = pool_max(m, __ldg(p + W + 1));
    y[idx] = m;
  }
}

// Backward: the whole window gradient goes to the FIRST maximal element in row-major window
// order (strict `>`; NaN wins) — ATen's max_pool2d index rule.  The argmax is
// recomputed from the saved input, so no index tensor is stored.  Threads cover ceil(H/2) x
// ceil(W/2) cells so the odd trailing row/column (outside every window) is zero-filled.
__global__ void maxpool2x2_bwd_kernel(const float* __restrict__ gy, const float* __restrict__ x,
                                      float* __restrict__ gx, int planes, int H, int W, int HP,
                                      int WP) {
  const int HC = (H + 1) / 2, WC = (W + 1) / 2;
  long long total = (long long)planes * HC * WC;
  long long stride = (long long)gridDim.x * blockDim.x;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total; idx += stride) {
    int cx = (int)(idx % WC);
    long long r = idx / WC;
    int cy = (int)(r % HC);
    long long pl = r / HC;
    long long base = (pl * H + 2 * cy) * W + 2 * cx;
    if (cy < HP && cx < WP) {
      float v[4] = {__ldg(x + base), __ldg(x + base + 1), __ldg(x + base + W), __ldg(x + base + W + 1)};
      float m = -INFINITY;
      int am = 0;
#pragma unroll
      for (int k = 0; k < 4; ++k)
        if (v[k] > m || v[k] != v[k]) { m = v[k]; am = k; }
      float g = __ldg(gy + (pl * HP + cy) * WP + cx);
      gx[base] = (am == 0) ? g : 0.f;
      gx[base + 1] = (am == 1) ? g : 0.f;
      gx[base + W] = (am == 2) ? g : 0.f;
      gx[base + W + 1] = (am == 3) ? g : 0.f;
    } else {
      for (int dy = 0; dy < 2; ++dy)
        for (int dx = 0; dx < 2; ++dx)
          if (2 * cy + dy < H && 2 * cx + dx < W) gx[base + dy * W + dx] = 0.f;
    }
  }
}

// ---- nearest upsample + concat (step1.py:78-90) ------------------------------------------
__global__ void upcat_fwd_kernel(const float* __restrict__ skip, const float* __restrict__ low,
                                 float* __restrict__ out, int B, int Cs, int Cl, int H, int W,
                                 int HL, int WL, int skip_first, NearestMap mh, NearestMap mw) {
  const int C = Cs + Cl;
  long long total = (long long)B * C * H * W;
  long long stride = (long long)gridDim.x * blockDim.x;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total; idx += stride) {
    int x = (int)(idx % W);
    long long r = idx / W;
    int y = (int)(r % H); r /= H;
    int ch = (int)(r % C);
    int b = (int)(r / C);
    bool is_skip = skip_first ? (ch < Cs) : (ch >= Cl);
    float v;
    if (is_skip) {
      int cs = skip_first ? ch : ch - Cl;
      v = __ldg(skip + (((long long)b * Cs + cs) * H + y) * W + x);
    } else {
      int cl = skip_first ? ch - Cs : ch;
      v = __ldg(low + (((long long)b * Cl + cl) * HL + mh(y)) * WL + mw(x));
    }
    out[idx] = v;
  }
}

__global__ void upcat_bwd_skip_kernel(const float* __restrict__ gout, float* __restrict__ gskip,
                                      int B, int Cs, int Cl, int H, int W, int skip_first) {
  long long total = (long long)B * Cs * H * W;
  long long stride = (long long)gridDim.x * blockDim.x;
  const int C = Cs + Cl;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total; idx += stride) {
    long long hw = idx % ((long long)H * W);
    long long r = idx / ((long long)H * W);
    int cs = (int)(r % Cs);
    int b = (int)(r / Cs);
    int ch = skip_first ? cs : Cl + cs;
    gskip[idx] = __ldg(gout + ((long long)b * C + ch) * H * W + hw);
  }
}

// glow[sy,sx] = sum of gout over every destination pixel whose nearest source is (sy,sx).
__global__ void upcat_bwd_low_kernel(const float* __restrict__ gout, float* __restrict__ glow,
                                     int B, int Cs, int Cl, int H, int W, int HL, int WL,
                                     int skip_first, NearestMap mh, NearestMap mw) {
  long long total = (long long)B * Cl * HL * WL;
  long long stride = (long long)gridDim.x * blockDim.x;
  const int C = Cs + Cl;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total; idx += stride) {
    int sx = (int)(idx % WL);
    long long r = idx / WL;
    int sy = (int)(r % HL); r /= HL;
    int cl = (int)(r % Cl);
    int b = (int)(r / Cl);
    int ch = skip_first ? Cs + cl : cl;
    const float* g = gout + ((long long)b * C + ch) * H * W;
    int y0, y1, x0, x1;
    if (mh.mode == 0) { y0 = sy; y1 = sy + 1; }
    else if (mh.mode == 1) { y0 = 2 * sy; y1 = 2 * sy + 2; }
    else { y0 = max(0, (int)floorf(sy / mh.scale) - 2); y1 = min(H, (int)ceilf((sy + 1) / mh.scale) + 2); }
    if (mw.mode == 0) { x0 = sx; x1 = sx + 1; }
    else if (mw.mode == 1) { x0 = 2 * sx; x1 = 2 * sx + 2; }
    else { x0 = max(0, (int)floorf(sx / mw.scale) - 2); x1 = min(W, (int)ceilf((sx + 1) / mw.scale) + 2); }
    float acc = 0.f;
    for (int y = y0; y < y1; ++y) {
      if (mh(y) != sy) continue;
      for (int x = x0; x < x1; ++x)
        if (mw(x) == sx) acc += __ldg(g + (long long)y * W + x);
    }
    glow[idx] = acc;
  }
}

// ---- bilinear, align_corners=True (step2.py:252,271) -------------------------------------
// Restates ATen's upsample_bilinear2d: scale=(in-1)/(out-1) in fp32, src=scale*dst,
// i0=min((int)src,in-1), i1=min(i0+1,in-1), l1=clamp(src-i0,0,1), l0=1-l1,
// out = h0*(w0*v00 + w1*v01) + h1*(w0*v10 + w1*v11).
__global__ void bilinear_ac_kernel(const float* __restrict__ x, float* __restrict__ y, int planes,
                                   int H, int W, int HO, int WO, float sh, float sw) {
  long long total = (long long)planes * HO * WO;
  long long stride = (long long)gridDim.x * blockDim.x;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total; idx += stride) {
    int ox = (int)(idx % WO);
    long long r = idx / WO;
    int oy = (int)(r % HO);
    long long pl = r / HO;
    float fy = sh * (float)oy, fx = sw * (float)ox;
    int y0 = min((int)fy, H - 1), x0 = min((int)fx, W - 1);
    int y1 = min(y0 + 1, H - 1), x1 = min(x0 + 1, W - 1);
    float ly1 = fminf(fmaxf(fy - (float)y0, 0.f), 1.f), lx1 = fminf(fmaxf(fx - (float)x0, 0.f), 1.f);
    float ly0 = 1.f - ly1, lx0 = 1.f - lx1;
    const float* p = x + pl * H * W;
    float v00 = __ldg(p + (long long)y0 * W + x0), v01 = __ldg(p + (long long)y0 * W + x1);
    float v10 = __ldg(p + (long long)y1 * W + x0), v11 = __ldg(p + (long long)y1 * W + x1);
    y[idx] = ly0 * (lx0 * v00 + lx1 * v01) + ly1 * (lx0 * v10 + lx1 * v11);
  }
}

__global__ void zero_border_kernel(float* __restrict__ x, int planes, int H, int W, int top,
                                   int bottom, int left) {
  long long total = (long long)planes * H * W;
  long long stride = (long long)gridDim.x * blockDim.x;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total; idx += stride) {
    int xx = (int)(idx % W);
    int yy = (int)((idx / W) % H);
    if (yy < top || yy >= H - bottom || xx < left) x[idx] = 0.f;
  }
}

__global__ void bn_fold_kernel(const float* gamma, const float* beta, const float* mean,
                               const float* var, const float* conv_bias, float* scale, float* shift,
                               int C, float eps) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= C) return;
  float sc = gamma[i] / sqrtf(var[i] + eps);
  float b = conv_bias ? conv_bias[i] : 0.f;
  scale[i] = sc;
  shift[i] = beta[i] + (b - mean[i]) * sc;
}

}  // namespace nconv

using namespace nconv;

extern "C" int nconv_b200_abi_version(void) { return 1; }

extern "C" unsigned long long nconv_b200_launch_count(void) {
  return g_launch_count.load(std::memory_order_relaxed);
}

extern "C" int nconv_b200_profile_enable(int on) {
  profile_clear();
  g_profile_on.store(on ? 1 : 0, std::memory_order_relaxed);
  return 0;
}

extern "C" int nconv_b200_profile_count(void) {
  std::lock_guard<std::mutex> lk(g_prof_mu);
  return (int)g_prof.size();
}

extern "C" int nconv_b200_profile_get(int i, char* label, int label_cap, float* ms, double* flops,
                                      double* bytes) {
  std::lock_guard<std::mutex> lk(g_prof_mu);
  if (i < 0 || i >= (int)g_prof.size() || !label || label_cap <= 0 || !ms || !flops || !bytes)
    return NCONV_ERR_BAD_ARG;
  ProfRecord* r = g_prof[i];
  snprintf(label, label_cap, "%s", r->label);
  *flops = r->flops; *bytes = r->bytes;
  return (int)cudaEventElapsedTime(ms, r->a, r->b);
}

extern "C" const char* nconv_b200_error_string(int code) {
  switch (code) {
    case 0: return "success";
    case NCONV_ERR_BAD_ARG: return "nconv_b200: bad argument (null pointer or non-positive size)";
    case NCONV_ERR_UNSUPPORTED: return "nconv_b200: unsupported shape/flag combination";
    case NCONV_ERR_WORKSPACE: return "nconv_b200: workspace too small";
    case NCONV_ERR_STATIC_SCRATCH: return "nconv_b200: size exceeds the static scratch of the no-allocation entry point (see its limit in include/nconv_b200.h); use the entry point that takes a workspace";
    default: return cudaGetErrorString((cudaError_t)code);
  }
}

extern "C" int nconv_b200_enforce_pos(float* w, long long n, int mode, int group, void* stream) {
  if (!w || n <= 0 || mode < 0 || mode > 3) return NCONV_ERR_BAD_ARG;
  if (mode == 3 && (group <= 0 || n % group != 0)) return NCONV_ERR_BAD_ARG;
  {
    LaunchScope ls((cudaStream_t)stream, "enforce_pos", 0.0, 0.0);
    enforce_pos_kernel<<<grid_for(mode == 3 ? n / group : n), 256, 0, (cudaStream_t)stream>>>(w, n, mode, group);
  }
  return (int)cudaGetLastError();
}

extern "C" int nconv_b200_confidence_mask(const float* s, float* c, long long n, float thr,
                                          void* stream) {
  if (!s || !c || n < 0) return NCONV_ERR_BAD_ARG;
  if (n == 0) return 0;
  {
    LaunchScope ls((cudaStream_t)stream, "confidence_mask", 0.0, 0.0);
    confidence_mask_kernel<<<grid_for(n), 256, 0, (cudaStream_t)stream>>>(s, c, n, thr);
  }
  return (int)cudaGetLastError();
}

extern "C" int nconv_b200_maxpool2x2_fwd(const float* x, float* y, int planes, int H, int W,
                                         void* stream) {
  if (!x || !y || planes <= 0 || H < 2 || W < 2) return NCONV_ERR_BAD_ARG;
  int HP = H / 2, WP = W / 2;
  {
    LaunchScope ls((cudaStream_t)stream, "maxpool2x2_fwd", 0.0, 0.0);
    maxpool2x2_fwd_kernel<<<grid_for((long long)planes * HP * WP), 256, 0, (cudaStream_t)stream>>>(
      x, y, planes, H, W, HP, WP);
  }
  return (int)cudaGetLastError();
}

extern "C" int nconv_b200_maxpool2x2_bwd(const float* gy, const float* x, float* gx, int planes,
                                         int H, int W, void* stream) {
  if (!gy || !x || !gx || planes <= 0 || H < 2 || W < 2) return NCONV_ERR_BAD_ARG;
  int HP = H / 2, WP = W / 2;
  long long cells = (long long)planes * ((H + 1) / 2) * ((W + 1) / 2);
  {
    LaunchScope ls((cudaStream_t)stream, "maxpool2x2_bwd", 0.0, 0.0);
    maxpool2x2_bwd_kernel<<<grid_for(cells), 256, 0, (cudaStream_t)stream>>>(gy, x, gx, planes, H, W, HP, WP);
  }
  return (int)cudaGetLastError();
}

// ---- opt-in "confidence-argmax" pooling: what upstream nconv (and BASELINE's north_star text) does and
// the reference does NOT (it pools data and confidence independently, step1.py:62-74): the confidence is max-pooled, the
// data is GATHERED at the confidence's argmax, and the pooled confidence is divided by 4.  Same window order and
// tie-break as F.max_pool2d(return_indices=True): first maximum in row-major order, strict `>`, NaN wins.
namespace nconv {
__global__ void pool_conf_argmax_fwd_kernel(const float* __restrict__ x, const float* __restrict__ c, float* __restrict__ xo,
                                            float* __restrict__ co, int planes, int H, int W, int HP, int WP) {
  const long long total = (long long)planes * HP * WP;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
    const int px = (int)(idx % WP);
    const long long r = idx / WP;
    const int py = (int)(r % HP);
    const long long base = ((r / HP) * H + 2 * py) * W + 2 * px;
    const int off[4] = {0, 1, W, W + 1};
    float m = -INFINITY;
    int am = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float v = __ldg(c + base + off[k]);
      if (v > m || v != v) { m = v; am = k; }
    }
    co[idx] = m * 0.25f;
    xo[idx] = __ldg(x + base + off[am]);
  }
}

// gx, gc <- scatter of (gxo, gco / 4) to the confidence's argmax; every other window element gets 0.
__global__ void pool_conf_argmax_bwd_kernel(const float* __restrict__ gxo, const float* __restrict__ gco, const float* __restrict__ c,
                                            float* __restrict__ gx, float* __restrict__ gc, int planes, int H, int W, int HP, int WP) {
  const int HC = (H + 1) / 2, WC = (W + 1) / 2;
  const long long total = (long long)planes * HC * WC;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
    const int cx = (int)(idx % WC);
    const long long r = idx / WC;
    const int cy = (int)(r % HC);
    const long long pl = r / HC;
    const long long base = (pl * H + 2 * cy) * W + 2 * cx;
    if (cy < HP && cx < WP) {
      const int off[4] = {0, 1, W, W + 1};
      float m = -INFINITY;
      int am = 0;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float v = __ldg(c + base + off[k]);
        if (v > m || v != v) { m = v; am = k; }
      }
      const long long o = (pl * HP + cy) * WP + cx;
      const float g1 = gxo ? __ldg(gxo + o) : 0.f, g2 = gco ? 0.25f * __ldg(gco + o) : 0.f;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        if (gx) gx[base + off[k]] = (am == k) ? g1 : 0.f;
        if (gc) gc[base + off[k]] = (am == k) ? g2 : 0.f;
      }
    } else {
      for (int dy = 0; dy < 2; ++dy)
        for (int dx = 0; dx < 2; ++dx)
          if (2 * cy + dy < H && 2 * cx + dx < W) {
            if (gx) gx[base + dy * W + dx] = 0.f;
            if (gc) gc[base + dy * W + dx] = 0.f;
          }
    }
  }
}
}  // namespace nconv

extern "C" int nconv_b200_pool_conf_argmax_fwd(const float* x, const float* c, float* xo, float* co, int planes, int H, int W,
                                               void* stream) {
  if (!x || !c || !xo || !co || planes <= 0 || H < 2 || W < 2) return NCONV_ERR_BAD_ARG;
  const int HP = H / 2, WP = W / 2;
  {
    LaunchScope ls((cudaStream_t)stream, "pool_conf_argmax_fwd", 0.0, 4.0 * planes * (2.0 * H * W + 2.0 * HP * WP));
    nconv::pool_conf_argmax_fwd_kernel<<<grid_for((long long)planes * HP * WP), 256, 0, (cudaStream_t)stream>>>(x, c, xo, co, planes, H, W, HP, WP);
  }
  return (int)cudaGetLastError();
}

extern "C" int nconv_b200_pool_conf_argmax_bwd(const float* gxo, const float* gco, const float* c, float* gx, float* gc, int planes,
                                               int H, int W, void* stream) {
  if (!c || (!gx && !gc) || planes <= 0 || H < 2 || W < 2) return NCONV_ERR_BAD_ARG;
  const int HP = H / 2, WP = W / 2;
  const long long cells = (long long)planes * ((H + 1) / 2) * ((W + 1) / 2);
  {
    LaunchScope ls((cudaStream_t)stream, "pool_conf_argmax_bwd", 0.0, 0.0);
    nconv::pool_conf_argmax_bwd_kernel<<<grid_for(cells), 256, 0, (cudaStream_t)stream>>>(gxo, gco, c, gx, gc, planes, H, W, HP, WP);
  }
  return (int)cudaGetLastError();
}

extern "C" int nconv_b200_upcat_fwd(const float* skip, const float* low, float* out, int B, int Cs,
                                    int Cl, int H, int W, int HL, int WL, int skip_first,
                                    void* stream) {
  if (!skip || !low || !out || B <= 0 || Cs <= 0 || Cl <= 0 || H <= 0 || W <= 0 || HL <= 0 || WL <= 0)
    return NCONV_ERR_BAD_ARG;
  long long total = (long long)B * (Cs + Cl) * H * W;
  {
    LaunchScope ls((cudaStream_t)stream, "upcat_fwd", 0.0, 0.0);
    upcat_fwd_kernel<<<grid_for(total), 256, 0, (cudaStream_t)stream>>>(
      skip, low, out, B, Cs, Cl, H, W, HL, WL, skip_first, NearestMap::make(HL, H), NearestMap::make(WL, W));
  }
  return (int)cudaGetLastError();
}

extern "C" int nconv_b200_upcat_bwd(const float* gout, float* gskip, float* glow, int B, int Cs,
                                    int Cl, int H, int W, int HL, int WL, int skip_first,
                                    void* stream) {
  if (!gout || B <= 0 || Cs <= 0 || Cl <= 0 || H <= 0 || W <= 0 || HL <= 0 || WL <= 0)
    return NCONV_ERR_BAD_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  if (gskip) {
    {
      LaunchScope ls(st, "upcat_bwd_skip", 0.0, 0.0);
      upcat_bwd_skip_kernel<<<grid_for((long long)B * Cs * H * W), 256, 0, st>>>(gout, gskip, B, Cs, Cl, H, W, skip_first);
    }
  }
  if (glow) {
    {
      LaunchScope ls(st, "upcat_bwd_low", 0.0, 0.0);
      upcat_bwd_low_kernel<<<grid_for((long long)B * Cl * HL * WL), 256, 0, st>>>(
        gout, glow, B, Cs, Cl, H, W, HL, WL, skip_first, NearestMap::make(HL, H), NearestMap::make(WL, W));
    }
  }
  return (int)cudaGetLastError();
}

extern "C" int nconv_b200_bilinear_ac_fwd(const float* x, float* y, int planes, int H, int W,
                                          int HO, int WO, void* stream) {
  if (!x || !y || planes <= 0 || H <= 0 || W <= 0 || HO <= 0 || WO <= 0) return NCONV_ERR_BAD_ARG;
  float sh = HO > 1 ? (float)(H - 1) / (float)(HO - 1) : 0.f;
  float sw = WO > 1 ? (float)(W - 1) / (float)(WO - 1) : 0.f;
  {
    LaunchScope ls((cudaStream_t)stream, "bilinear_ac", 0.0, 0.0);
    bilinear_ac_kernel<<<grid_for((long long)planes * HO * WO), 256, 0, (cudaStream_t)stream>>>(
      x, y, planes, H, W, HO, WO, sh, sw);
  }
  return (int)cudaGetLastError();
}

extern "C" int nconv_b200_zero_border(float* x, int planes, int H, int W, int top, int bottom,
                                      int left, void* stream) {
  if (!x || planes <= 0 || H <= 0 || W <= 0 || top < 0 || bottom < 0 || left < 0) return NCONV_ERR_BAD_ARG;
  {
    LaunchScope ls((cudaStream_t)stream, "zero_border", 0.0, 0.0);
    zero_border_kernel<<<grid_for((long long)planes * H * W), 256, 0, (cudaStream_t)stream>>>(
      x, planes, H, W, top, bottom, left);
  }
  return (int)cudaGetLastError();
}

extern "C" int nconv_b200_bn_fold(const float* gamma, const float* beta, const float* mean,
                                  const float* var, const float* conv_bias, float* scale,
                                  float* shift, int C, float eps, void* stream) {
  if (!gamma || !beta || !mean || !var || !scale || !shift || C <= 0) return NCONV_ERR_BAD_ARG;
  {
    LaunchScope ls((cudaStream_t)stream, "bn_fold", 0.0, 0.0);
    bn_fold_kernel<<<(C + 127) / 128, 128, 0, (cudaStream_t)stream>>>(gamma, beta, mean, var, conv_bias,
                                                                  scale, shift, C, eps);
  }
  return (int)cudaGetLastError();
}
