// Whole DNET.forward (inference) as one enqueue of 9 fused kernels —
// /root/reference/models/step1.py:51-94.  The reference issues ~100 eager ATen kernels for the
// same graph; here every max-pool is an extra epilogue output of the producing
// NConv kernel, every nearest-upsample + concat is a load-side index map of the consuming one,
// the threshold is the prologue of nconv1 and nconv6 -> nconv7 -> crop is a single kernel.
#include "common.cuh"
#include "nconv_b200.h"
#include "nconv_tc.h"
#include <cstdio>
#include <cstdlib>

namespace {

struct Level { int H, W; };

struct DnetPlan {
  Level l[4];
  size_t off[32];
  size_t total_floats;
};

// Buffer ids (each holds [n,8,H_l,W_l] floats)
enum {
  X1A, C1A, X1, C1, X1P, C1P, X2, C2, X2P, C2P, X3, C3, X3P, C3P, X4, C4, X34, C34, X23, C23, NBUF
};
static const int kBufLevel[NBUF] = {0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 2, 2, 1, 1};

static DnetPlan make_plan(int n, int H, int W) {
  DnetPlan p;
  p.l[0] = {H, W};
  for (int i = 1; i < 4; ++i) p.l[i] = {p.l[i - 1].H / 2, p.l[i - 1].W / 2};
  size_t cur = 0;
  for (int i = 0; i < NBUF; ++i) {
    p.off[i] = cur;
    size_t fl = (size_t)n * 8 * p.l[kBufLevel[i]].H * p.l[kBufLevel[i]].W;
    cur += (fl + 63) & ~(size_t)63;  // keep every buffer 256-byte aligned
  }
  p.total_floats = cur;
  return p;
}

// ---- tensor-core pipeline (csrc/nconv_tc.cu): every activation in the N8 split-bf16 layout, the seven 8-channel
// layers on tcgen05, nconv1 on CUDA cores writing N8.  Needs exact 2x levels (H, W multiples of 8: pools floor and
// the nearest upsample is "repeat 2x2") and a lowest level at least one staged row wide.
enum { A0, TX1, TP1, TX2, TP2, TX3, TP3, TU4, TU34, TU23, TNBUF };
static const int kTcLevel[TNBUF] = {0, 0, 1, 1, 2, 2, 3, 2, 1, 0};

struct DnetTcPlan {
  Level l[4];
  size_t off[TNBUF], woff[7];
  size_t total_bytes;
};

static bool tc_eligible(int H, int W) {
  static const bool off = getenv("NCONV_B200_NO_NCONV_TC") != nullptr;      // A/B measurements only
  return !off && H % 8 == 0 && W % 8 == 0 && W / 8 + 4 >= 32 && H / 8 >= 4;
}

static DnetTcPlan make_tc_plan(int n, int H, int W) {
  DnetTcPlan p;
  p.l[0] = {H, W};
  for (int i = 1; i < 4; ++i) p.l[i] = {p.l[i - 1].H / 2, p.l[i - 1].W / 2};
  size_t cur = 0;
  for (int i = 0; i < TNBUF; ++i) {
    p.off[i] = cur;
    cur += (nconv_b200_n8_bytes(n, p.l[kTcLevel[i]].H, p.l[kTcLevel[i]].W) + 255) & ~(size_t)255;
  }
  for (int i = 0; i < 7; ++i) {
    p.woff[i] = cur;
    cur += (nconv_b200_nconv_tc_wpack_bytes(i < 4 ? 5 : 3, i < 4 ? 0 : 8) + 255) & ~(size_t)255;
  }
  p.total_bytes = cur;
  return p;
}

static int dnet_forward_tc(const float* S, const float* const* params, float* out, int B, int H, int W, void* workspace,
                           void* stream) {
  const DnetTcPlan p = make_tc_plan(B, H, W);
  uint8_t* ws = (uint8_t*)workspace;
  auto buf = [&](int id) { return (void*)(ws + p.off[id]); };
  cudaStream_t st = (cudaStream_t)stream;
  int rc;
  {
    nconv::N8RingJobs rj = {};
    rj.B = B; rj.n = TNBUF;
    for (int i = 0; i < TNBUF; ++i) { rj.buf[i] = buf(i); rj.H[i] = p.l[kTcLevel[i]].H; rj.W[i] = p.l[kTcLevel[i]].W; }
    if ((rc = nconv::n8_zero_rings_multi(rj, st))) return rc;
    // layers 1..7 of the parameter list: nconv2, nconv_down1..3 (8->8 5x5), nconv4, nconv5 (skip first), nconv6 (upsampled first)
    nconv::NtPackJobs pj = {};
    pj.n = 7;
    for (int i = 0; i < 7; ++i) {
      pj.w[i] = params[2 * (i + 1)]; pj.wpack[i] = ws + p.woff[i];
      pj.K[i] = i < 4 ? 5 : 3; pj.nsrc[i] = i < 4 ? 1 : 2; pj.a_first[i] = (i == 6) ? 0 : 1;
    }
    if ((rc = nconv::nconv_tc_pack_multi(pj, st))) return rc;
  }
  const float eps = 1e-7f;
  // NCONV_B200_DNET_SYNC=1 (debugging only): synchronise after every launch and name the one that failed
  static const bool dbg_sync = getenv("NCONV_B200_DNET_SYNC") != nullptr;
  auto check = [&](const char* what) {
    if (!dbg_sync) return 0;
    const cudaError_t e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) fprintf(stderr, "[nconv_b200] dnet_forward_tc: %s failed: %s\n", what, cudaGetErrorString(e));
    return (int)e;
  };
  if ((rc = check("prep"))) return rc;
  if ((rc = nconv_b200_nconv1_n8(S, params[0], params[1], buf(A0), B, H, W, 0.01f, eps, stream))) return rc;   // step1.py:53-56
  if ((rc = check("nconv1"))) return rc;
  auto layer = [&](int i, int lvl) {
    nconv_b200_nconv_tc_desc d = {};
    d.wpack = ws + p.woff[i]; d.w = params[2 * (i + 1)]; d.bias = params[2 * (i + 1) + 1];
    d.B = B; d.H = p.l[lvl].H; d.W = p.l[lvl].W; d.K = i < 4 ? 5 : 3; d.CB = i < 4 ? 0 : 8; d.eps = eps;
    return d;
  };
  { nconv_b200_nconv_tc_desc d = layer(0, 0); d.a = buf(A0); d.out_n8 = buf(TX1); d.out_pool = buf(TP1);       // nconv2 + pools, :57-64
    if ((rc = nconv_b200_nconv_tc(&d, stream))) return rc;
    if ((rc = check("nconv2"))) return rc; }
  { nconv_b200_nconv_tc_desc d = layer(1, 1); d.a = buf(TP1); d.out_n8 = buf(TX2); d.out_pool = buf(TP2);      // nconv_down1 + pools, :65-69
    if ((rc = nconv_b200_nconv_tc(&d, stream))) return rc;
    if ((rc = check("nconv_down1"))) return rc; }
  { nconv_b200_nconv_tc_desc d = layer(2, 2); d.a = buf(TP2); d.out_n8 = buf(TX3); d.out_pool = buf(TP3);      // nconv_down2 + pools, :70-74
    if ((rc = nconv_b200_nconv_tc(&d, stream))) return rc;
    if ((rc = check("nconv_down2"))) return rc; }
  { nconv_b200_nconv_tc_desc d = layer(3, 3); d.a = buf(TP3); d.out_up = buf(TU4);                             // nconv_down3, upsampled for :78-79
    if ((rc = nconv_b200_nconv_tc(&d, stream))) return rc;
    if ((rc = check("nconv_down3"))) return rc; }
  { nconv_b200_nconv_tc_desc d = layer(4, 2); d.a = buf(TX3); d.b = buf(TU4); d.out_up = buf(TU34);            // nconv4 on cat(x3, up(x4)), :80, :83-84
    if ((rc = nconv_b200_nconv_tc(&d, stream))) return rc;
    if ((rc = check("nconv4"))) return rc; }
  { nconv_b200_nconv_tc_desc d = layer(5, 1); d.a = buf(TX2); d.b = buf(TU34); d.out_up = buf(TU23);           // nconv5 on cat(x2, up(x34)), :85, :88-89
    if ((rc = nconv_b200_nconv_tc(&d, stream))) return rc;
    if ((rc = check("nconv5"))) return rc; }
  { nconv_b200_nconv_tc_desc d = layer(6, 0); d.a = buf(TX1); d.b = buf(TU23);                                  // nconv6 on cat(up(x23), x1) -> nconv7 -> crop, :90-94
    d.w_tail = params[16]; d.b_tail = params[17]; d.out_tail = out;
    if ((rc = nconv_b200_nconv_tc(&d, stream))) return rc;
    if ((rc = check("nconv6"))) return rc; }
  return 0;
}

}  // namespace

extern "C" size_t nconv_b200_dnet_workspace_bytes(int B, int H, int W) {
  if (B <= 0 || H < 8 || W < 8) return 0;
  size_t n = make_plan(B, H, W).total_floats * sizeof(float);
  if (tc_eligible(H, W)) {
    const size_t t = make_tc_plan(B, H, W).total_bytes;
    if (t > n) n = t;
  }
  return n;
}

static int dnet_forward_impl(const float* S, const float* const* params, float* out, int B, int H, int W, void* workspace,
                             size_t workspace_bytes, void* stream, bool allow_tc) {
  if (!S || !params || !out || !workspace || B <= 0) return NCONV_ERR_BAD_ARG;
  if (H < 8 || W < 8) return NCONV_ERR_BAD_ARG;  // three 2x2 pools must leave >= 1 pixel
  for (int i = 0; i < 18; ++i)
    if (!params[i]) return NCONV_ERR_BAD_ARG;
  if (workspace_bytes < nconv_b200_dnet_workspace_bytes(B, H, W)) return NCONV_ERR_WORKSPACE;
  if (allow_tc && tc_eligible(H, W) && (reinterpret_cast<uintptr_t>(workspace) & 255u) == 0)
    return dnet_forward_tc(S, params, out, B, H, W, workspace, stream);
  DnetPlan p = make_plan(B, H, W);
  float* ws = (float*)workspace;
  auto buf = [&](int id) { return ws + p.off[id]; };
  const float eps = 1e-7f;  // step1.py:103

  auto base = [&](int layer, int lvl) {
    nconv_b200_fwd_desc d = {};
    d.w = params[2 * layer]; d.bias = params[2 * layer + 1];
    d.B = B; d.CO = 8; d.H = p.l[lvl].H; d.W = p.l[lvl].W;
    d.a_first = 1; d.eps = eps; d.thr = 0.01f;  // step1.py:53
    return d;
  };
  int rc;
  {  // nconv1: threshold prologue, 1->8, 5x5 pad 2 (step1.py:53-56)
    nconv_b200_fwd_desc d = base(0, 0);
    d.xa = S; d.CA = 1; d.K = 5; d.pad = 2; d.flags = NCONV_B200_F_THRESH;
    d.y = buf(X1A); d.cout = buf(C1A);
    if ((rc = nconv_b200_nconv2d_fwd_fused(&d, stream))) return rc;
  }
  {  // nconv2 + pools (step1.py:57-64)
    nconv_b200_fwd_desc d = base(1, 0);
    d.xa = buf(X1A); d.ca = buf(C1A); d.CA = 8; d.K = 5; d.pad = 2; d.flags = NCONV_B200_F_POOL;
    d.y = buf(X1); d.cout = buf(C1); d.y_pool = buf(X1P); d.c_pool = buf(C1P);
    if ((rc = nconv_b200_nconv2d_fwd_fused(&d, stream))) return rc;
  }
  {  // nconv_down1 + pools (step1.py:65-69)
    nconv_b200_fwd_desc d = base(2, 1);
    d.xa = buf(X1P); d.ca = buf(C1P); d.CA = 8; d.K = 5; d.pad = 2; d.flags = NCONV_B200_F_POOL;
    d.y = buf(X2); d.cout = buf(C2); d.y_pool = buf(X2P); d.c_pool = buf(C2P);
    if ((rc = nconv_b200_nconv2d_fwd_fused(&d, stream))) return rc;
  }
  {  // nconv_down2 + pools (step1.py:70-74)
    nconv_b200_fwd_desc d = base(3, 2);
    d.xa = buf(X2P); d.ca = buf(C2P); d.CA = 8; d.K = 5; d.pad = 2; d.flags = NCONV_B200_F_POOL;
    d.y = buf(X3); d.cout = buf(C3); d.y_pool = buf(X3P); d.c_pool = buf(C3P);
    if ((rc = nconv_b200_nconv2d_fwd_fused(&d, stream))) return rc;
  }
  {  // nconv_down3 (step1.py:75)
    nconv_b200_fwd_desc d = base(4, 3);
    d.xa = buf(X3P); d.ca = buf(C3P); d.CA = 8; d.K = 5; d.pad = 2; d.flags = 0;
    d.y = buf(X4); d.cout = buf(C4);
    if ((rc = nconv_b200_nconv2d_fwd_fused(&d, stream))) return rc;
  }
  {  // nconv4 on cat(skip x3, up(x4)) (step1.py:78-80)
    nconv_b200_fwd_desc d = base(5, 2);
    d.xa = buf(X3); d.ca = buf(C3); d.CA = 8; d.xb = buf(X4); d.cb = buf(C4); d.CB = 8;
    d.HB = p.l[3].H; d.WB = p.l[3].W; d.K = 3; d.pad = 1; d.a_first = 1;
    d.y = buf(X34); d.cout = buf(C34);
    if ((rc = nconv_b200_nconv2d_fwd_fused(&d, stream))) return rc;
  }
  {  // nconv5 on cat(skip x2, up(x34)) (step1.py:83-85)
    nconv_b200_fwd_desc d = base(6, 1);
    d.xa = buf(X2); d.ca = buf(C2); d.CA = 8; d.xb = buf(X34); d.cb = buf(C34); d.CB = 8;
    d.HB = p.l[2].H; d.WB = p.l[2].W; d.K = 3; d.pad = 1; d.a_first = 1;
    d.y = buf(X23); d.cout = buf(C23);
    if ((rc = nconv_b200_nconv2d_fwd_fused(&d, stream))) return rc;
  }
  {  // nconv6 on cat(up(x23), skip x1) -> nconv7 -> crop (step1.py:88-94)
    nconv_b200_fwd_desc d = base(7, 0);
    d.xa = buf(X1); d.ca = buf(C1); d.CA = 8; d.xb = buf(X23); d.cb = buf(C23); d.CB = 8;
    d.HB = p.l[1].H; d.WB = p.l[1].W; d.K = 3; d.pad = 0; d.a_first = 0;
    d.flags = NCONV_B200_F_TAIL;
    d.w_tail = params[16]; d.b_tail = params[17]; d.out_tail = out;
    if ((rc = nconv_b200_nconv2d_fwd_fused(&d, stream))) return rc;
  }
  return 0;
}

extern "C" int nconv_b200_dnet_forward(const float* S, const float* const* params, float* out, int B, int H, int W,
                                       void* workspace, size_t workspace_bytes, void* stream) {
  return dnet_forward_impl(S, params, out, B, H, W, workspace, workspace_bytes, stream, true);
}

extern "C" int nconv_b200_dnet_forward_fp32(const float* S, const float* const* params, float* out, int B, int H, int W,
                                            void* workspace, size_t workspace_bytes, void* stream) {
  return dnet_forward_impl(S, params, out, B, H, W, workspace, workspace_bytes, stream, false);
}
