/* nconv_b200 — C ABI of the B200 (sm_100a) normalized-convolution hot path.
 *
 * The reference (lllllcf/Realtime-Depth-Estimation-Nconv) is pure Python/PyTorch and has no
 * FFI of its own; its "operator API" for this path is nn.Module.forward.  Each entry point
 * below cites the reference lines whose ATen call sequence it replaces.  A maintainer binds
 * them with ctypes; `realtime-depth-estimation-nconv_b200/nconv_b200/`
 * is that binding, registered with torch.library.
 *
 * Conventions
 *   - all tensors: contiguous fp32 NCHW device memory owned by the caller
 *   - `stream` is a cudaStream_t passed as void*; work is enqueued, never synchronised
 *   - no allocation inside; scratch is passed in (`workspace`, size from *_workspace_bytes)
 *   - return value: 0, or a cudaError_t, or an NCONV_B200_ERR_* code; never throws
 *   - thread-safe for concurrent callers on different devices (nn.DataParallel style)
 */
#ifndef NCONV_B200_H_
#define NCONV_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NCONV_B200_ERR_BAD_ARG 100001
#define NCONV_B200_ERR_UNSUPPORTED 100002
#define NCONV_B200_ERR_WORKSPACE 100003
#define NCONV_B200_ERR_STATIC_SCRATCH 100004 /* size beyond the static scratch of a *_noalloc entry point */

/* flags of nconv_b200_fwd_desc */
#define NCONV_B200_F_THRESH 1 /* conf = (xa > thr), xa used raw  (step1.py:53-54)            */
#define NCONV_B200_F_POOL 2   /* also emit 2x2/2 max-pools of y and cout (step1.py:62-74)     */
#define NCONV_B200_F_TAIL 4   /* nconv6(valid 3x3) -> nconv7(1x1,pad 2) -> crop (step1.py:90-94) */

int nconv_b200_abi_version(void);
const char* nconv_b200_error_string(int code);
/* number of kernels this library has enqueued in the calling process (all devices) */
unsigned long long nconv_b200_launch_count(void);

/* Optional per-launch timing for the bench's roofline leg (no reference counterpart): when
 * enabled every launcher brackets its kernel with two CUDA events on the launching stream and
 * records a label and the launch's algorithmic flops/bytes.  enable(on) also clears the records.
 * profile_get returns the cudaEventElapsedTime status (synchronise the stream first). */
int nconv_b200_profile_enable(int on);
int nconv_b200_profile_count(void);
int nconv_b200_profile_get(int i, char* label, int label_cap, float* ms, double* flops,
                           double* bytes);

/* ---- (a3) NConv2d.forward — /root/reference/models/step1.py:116-149 -------------------
 * D = conv(c,W); N = conv(x*c,W); y = N/(D+eps)+b; cout = D/sum(W[o]).  stride 1, dil 1,
 * groups 1, zero padding (ph,pw).  One fused kernel. */
int nconv_b200_nconv2d_fwd(const float* x, const float* c, const float* w, const float* bias,
                           float* y, float* cout, int B, int Ci, int Co, int H, int W, int kh,
                           int kw, int ph, int pw, float eps, void* stream);

/* ---- fused stages of DNET.forward — step1.py:51-94 -------------------------------------
 * One NConv layer whose input is [source A | nearest-upsampled source B] channel-concatenated
 * (a_first: A's channels come first in the weight's input-channel axis — step1.py:80,85 —
 * else B first — step1.py:90), optionally thresholding A into a confidence (step1.py:53),
 * optionally also emitting the two independent 2x2 max-pools (step1.py:62-74), optionally
 * finishing with the nconv7 1x1 + crop tail (step1.py:92-94). */
typedef struct {
  const float* xa;  /* [B,CA,H,W] data of source A                                   */
  const float* ca;  /* [B,CA,H,W] confidence of A (ignored with F_THRESH)            */
  const float* xb;  /* [B,CB,HB,WB] data of source B (NULL when CB==0)               */
  const float* cb;  /* [B,CB,HB,WB]                                                  */
  const float* w;   /* [CO,CA+CB,K,K]                                                */
  const float* bias;/* [CO]                                                          */
  float* y;         /* [B,CO,HO,WO], HO=H+2*pad-K+1; may be NULL with F_TAIL         */
  float* cout;      /* [B,CO,HO,WO]                                                  */
  float* y_pool;    /* [B,CO,HO/2,WO/2] with F_POOL                                  */
  float* c_pool;
  const float* w_tail; /* [1,CO,1,1] nconv7 weight, with F_TAIL                      */
  const float* b_tail; /* [1]                                                        */
  float* out_tail;     /* [B,1,H,W]                                                  */
  int B, CA, CB, CO, H, W, HB, WB, K, pad;
  int a_first;
  int flags;
  float eps, thr;
} nconv_b200_fwd_desc;
int nconv_b200_nconv2d_fwd_fused(const nconv_b200_fwd_desc* d, void* stream);

/* ---- (a3, backward) analytic gradient of NConv2d.forward ----------------
 * Consumes the saved forward outputs (y, cout) instead of recomputing D and N:
 * D = cout*s, N/(D+eps) = y-b.  gx/gc: [B,Ci,H,W]; gw: [Co,Ci,kh,kw]; gb: [Co].
 * gx or gc may be NULL (input needs no grad).  workspace: nconv_b200_nconv2d_bwd_workspace_bytes. */
size_t nconv_b200_nconv2d_bwd_workspace_bytes(int B, int Ci, int Co, int H, int W, int kh, int kw,
                                              int ph, int pw);
int nconv_b200_nconv2d_bwd(const float* gy, const float* gcout, const float* x, const float* c,
                           const float* w, const float* bias, const float* y, const float* cout,
                           float* gx, float* gc, float* gw, float* gb, int B, int Ci, int Co, int H,
                           int W, int kh, int kw, int ph, int pw, float eps, void* workspace,
                           size_t workspace_bytes, void* stream);

/* ---- (a2) EnforcePos — step1.py:190-212: w <- pos_fn(w) in place (training only) --------
 * mode 0 softplus(beta=10), 1 exp, 2 sigmoid, 3 softmax over each [kh*kw] group (group=kh*kw). */
int nconv_b200_enforce_pos(float* w, long long n, int mode, int group, void* stream);

/* ---- (a4) threshold — step1.py:53: c0 = (S > thr) as fp32 ------------------------------ */
int nconv_b200_confidence_mask(const float* s, float* c, long long n, float thr, void* stream);

/* ---- (a5) 2x2/2 max-pool — step1.py:62-74 (F.max_pool2d(x,2,2)); first-max tie-break ---- */
int nconv_b200_maxpool2x2_fwd(const float* x, float* y, int planes, int H, int W, void* stream);
int nconv_b200_maxpool2x2_bwd(const float* gy, const float* x, float* gx, int planes, int H,
                              int W, void* stream);

/* ---- opt-in confidence-argmax pooling (NOT what the reference does, see step1.py:62-74) ----
 * co = maxpool2x2(c) / 4, xo = x gathered at c's argmax (first maximum in row-major window order, strict >,
 * NaN wins: F.max_pool2d(return_indices=True)).  bwd scatters (gxo, gco/4) to that argmax. */
int nconv_b200_pool_conf_argmax_fwd(const float* x, const float* c, float* xo, float* co, int planes,
                                    int H, int W, void* stream);
int nconv_b200_pool_conf_argmax_bwd(const float* gxo, const float* gco, const float* c, float* gx,
                                    float* gc, int planes, int H, int W, void* stream);

/* ---- (a6) nearest upsample + channel concat — step1.py:78-90 --------------------------- */
int nconv_b200_upcat_fwd(const float* skip, const float* low, float* out, int B, int Cs, int Cl,
                         int H, int W, int HL, int WL, int skip_first, void* stream);
int nconv_b200_upcat_bwd(const float* gout, float* gskip, float* glow, int B, int Cs, int Cl,
                         int H, int W, int HL, int WL, int skip_first, void* stream);

/* ---- (a4-a8) whole DNET.forward, inference — step1.py:51-94 ----------------------------
 * params: 18 pointers {w,b} for nconv1,nconv2,nconv_down1..3,nconv4,nconv5,nconv6,nconv7.
 * S: [B,1,H,W] -> out: [B,1,H,W] (crop generalised to [1:H+1,1:W+1]).  The last layer's confidence
 * is dropped as in the reference (step1.py:92-94); the opt-in (depth, confidence) return goes
 * through the staged ops (models/step1.py::SETP1_NCONV.forward_with_confidence). */
size_t nconv_b200_dnet_workspace_bytes(int B, int H, int W);
int nconv_b200_dnet_forward(const float* S, const float* const* params, float* out, int B, int H,
                            int W, void* workspace, size_t workspace_bytes, void* stream);
/* nconv_b200_dnet_forward runs the seven 8-channel layers on tcgen05 (split-bf16 operands, fp32 accumulate: y and cout
 * within ~5e-6 of fp32) when H and W are multiples of 8 and W >= 224, else the fp32 CUDA-core pipeline; _fp32 always
 * runs the latter (the frozen step 1 inside a training step uses it, so the training-step numerics stay pure fp32). */
int nconv_b200_dnet_forward_fp32(const float* S, const float* const* params, float* out, int B,
                                 int H, int W, void* workspace, size_t workspace_bytes, void* stream);

/* Weight gradient of a 3x3 / stride-1 / pad-1 convolution on tcgen05 (the dL/dW half of train_step2.py:60-66's
 * backward for step2.py:178-195, 216-236, 290-297): x and g = dL/dy are P8 tensors [B,CI,H,W] / [B,CO,H,W]
 * (nconv_b200_p8_pack), gw is fp32 [CO][CI][3][3].  CI in {32, 64, 128, 192, 256}, CO in {32, 64}.  Split-bf16
 * operands with all four partial products, fp32 accumulation in TMEM, deterministic final sum. */
size_t nconv_b200_conv_wgrad_tc_ws_bytes(int B, int CI, int CO, int H, int W);
int nconv_b200_conv_wgrad_tc(const void* x_p8, const void* g_p8, float* gw, void* workspace,
                             size_t workspace_bytes, int B, int CI, int CO, int H, int W, void* stream);

/* ---- GPU input pipeline: the loaders' per-sample host work after the H2D copy ----------------
 * rgb:  uint8 [B,H,W,3] (cv2.imread BGR, HWC) -> fp32 [B,3,H,W]           dataset/nyuloader.py:69, kittiloader.py:63
 * depth: [n] uint8 / uint16 / fp32 (elem_bytes 1/2/4) -> fp32 * scale      kittiloader.py:66-76 (`/ 256.0`)
 * sparsify: out = raw * (1 + noise) * mask, noise (fp32) and mask (uint8 0/1) optional   nyuloader.py:83-124 */
int nconv_b200_rgb_u8_to_f32(const uint8_t* hwc, float* chw, int B, int H, int W, void* stream);
int nconv_b200_depth_decode(const void* raw, int elem_bytes, float* out, long long n, float scale, void* stream);
int nconv_b200_sparsify(const float* raw, const float* noise, const uint8_t* mask, float* out, long long n, void* stream);

/* Debug aid (no reference counterpart): 64 bytes of pinned, device-mapped host memory in which a tcgen05 NConv kernel
 * records {barrier code, block, warp, tile / M-tile index, parity} when one of its bounded barrier waits gives up. */
int nconv_b200_debug_buffer(void* pinned_host_64_bytes);

/* ---- tensor-core NConv2d for DNET's FMA-bound layers (8->8 5x5, [8|8]->8 3x3) — step1.py:116-149 on tcgen05 ----
 * "N8" activation layout: an 8-channel (data, confidence) pair held as the two streams the layer
 * correlates, split to bf16 hi/lo (16 mantissa bits) with a 2-pixel zero ring:
 *     n8[n][stream: data*conf | conf][part: hi | lo][H+4][W+4][8]      (bf16, 64 B per pixel)
 * Same bytes as the fp32 (data, conf) pair.  Kernels write interior pixels only; the ring is
 * established by nconv_b200_n8_zero_ring.  pack/unpack convert from/to fp32 NCHW (tests, callers). */
size_t nconv_b200_n8_bytes(int B, int H, int W);
int nconv_b200_n8_zero_ring(void* n8, int B, int H, int W, void* stream);
int nconv_b200_n8_pack(const float* x, const float* c, void* n8, int B, int H, int W, void* stream);
int nconv_b200_n8_unpack(const void* n8, float* xc, float* c, int B, int H, int W, void* stream);

/* nconv1 of DNET (step1.py:53-56): c0 = (S > thr), NConv2d 1->8 5x5 pad 2, output written as N8. */
int nconv_b200_nconv1_n8(const float* S, const float* w, const float* bias, void* out_n8, int B,
                         int H, int W, float thr, float eps, void* stream);

/* Packed split-bf16 weights of one tensor-core NConv layer: w [8][CA+CB][K][K], K in {3,5}; source A
 * holds weight input channels [0,8) when a_first (or CB == 0), else [8,16). */
size_t nconv_b200_nconv_tc_wpack_bytes(int K, int CB);
int nconv_b200_nconv_tc_pack_weights(const float* w, void* wpack, int K, int CB, int a_first,
                                     void* stream);

/* One NConv2d layer (8 or 8+8 -> 8 channels, K in {3,5}, pad K/2, stride 1) on N8 inputs of the SAME
 * resolution (a nearest-upsampled source is supplied already expanded: see out_up).  Outputs, each
 * optional (NULL = not written):
 *   out_n8    (y, cout) as N8 at H x W                      [skip connection / next layer]
 *   out_pool  N8 at H/2 x W/2 of the two INDEPENDENT 2x2 max-pools of y and cout (step1.py:62-74)
 *   out_up    N8 at 2H x 2W, every pixel repeated 2x2 (F.interpolate nearest, exact 2x; :78-89)
 *   y, cout   fp32 NCHW [B,8,H,W]
 *   out_tail  fp32 [B,1,H,W]: the layer evaluated as step1.py:90 (valid 3x3, here centred) followed
 *             by nconv7 (1x1, pad 2) and the crop of :92-94 (needs w_tail [8], b_tail [1]). */
typedef struct {
  const void* a; const void* b;          /* N8 sources (b may be NULL) */
  const void* wpack; const float* w; const float* bias;   /* packed weights, original fp32 weights (for sum W), bias */
  void* out_n8; void* out_pool; void* out_up; float* y; float* cout;
  const float* w_tail; const float* b_tail; float* out_tail;
  int B, H, W, K, CB;
  float eps;
} nconv_b200_nconv_tc_desc;
int nconv_b200_nconv_tc(const nconv_b200_nconv_tc_desc* d, void* stream);

/* ---- SETP2 building blocks — /root/reference/models/step2.py ---------------------------
 * Direct convolution y = act(scale[o]*conv(cat[a,b]) + shift[o]) + residual, covering
 * ConvBlock (:290-297), RGBEncoder (:134-154, BN folded to scale/shift), Basic2d (:178-195),
 * Conv1x1 (:130), Conv3x3 head + residual (:156, :255-257).  K in {1,3}, stride in {1,2},
 * pad = K/2.  scale/shift/residual/b may be NULL. */
typedef struct {
  const float* a;       /* [B,CA,H,W]                                            */
  const float* b;       /* [B,CB,H,W] concatenated after a, or NULL               */
  const float* w;       /* [CO,CA+CB,K,K]                                        */
  const float* scale;   /* [CO] or NULL (=1)                                     */
  const float* shift;   /* [CO] or NULL (=0)                                     */
  const float* residual;/* [B,CO,HO,WO] added after the activation, or NULL      */
  float* y;             /* [B,CO,HO,WO]                                          */
  int B, CA, CB, CO, H, W, K, stride, relu;
} nconv_b200_conv_desc;
int nconv_b200_conv2d_fwd(const nconv_b200_conv_desc* d, void* stream);

/* ConvTranspose2d(k=4,s=2,p=1,bias=False) of cat[a(1ch... CA), b(CB)] + scale/shift + ReLU —
 * Basic2dTrans, step2.py:197-214; w: [CA+CB, CO, 4, 4]. */
int nconv_b200_convT4x4s2_fwd(const float* a, const float* b, const float* w, const float* scale,
                              const float* shift, float* y, int B, int CA, int CB, int CO, int H,
                              int W, int relu, void* stream);

/* ---- training mode of the RGB-guided stack (train_step2.py:60-66 runs SETP2_BP_TRAIN under
 * model.train(): BatchNorm2d with batch statistics, autograd through every conv of step2.py) ---- */

/* scratch for the per-channel fp64 reductions below */
size_t nconv_b200_train_ws_bytes(int C);

/* nn.BatchNorm2d training forward (step2.py:141,188,212) fused with ReLU and the residual that
 * RGBEncoder adds after the activation (:151-152): y = act(gamma*(x-mean_B)/sqrt(var_B+eps)+beta)
 * (+residual); running_mean/var updated with `momentum` (unbiased variance) unless NULL;
 * scale/shift/save_mean/save_rstd: [C] outputs kept for backward. */
int nconv_b200_bn_train_fwd(const float* x, const float* gamma, const float* beta, float* running_mean,
                            float* running_var, const float* residual, float* y, float* scale,
                            float* shift, float* save_mean, float* save_rstd, void* ws, int B, int C,
                            int H, int W, float eps, float momentum, int relu, void* stream);

/* backward of the above: gz (grad of the BatchNorm input z), ggamma, gbeta from gy (grad of the
 * activation output), z and the saved scale/shift/mean/rstd. */
int nconv_b200_bn_train_bwd(const float* gy, const float* z, const float* scale, const float* shift,
                            const float* mean, const float* rstd, const float* gamma, float* gz,
                            float* ggamma, float* gbeta, void* ws, int B, int C, int H, int W, int relu,
                            void* stream);

/* backward of ReLU(conv + bias) (ConvBlock, step2.py:290-297): g = gy*[y>0] (g may be NULL),
 * gbias[c] = sum g; relu=0 gives the plain per-channel sum of gy. */
int nconv_b200_relu_bias_bwd(const float* gy, const float* y, float* g, float* gbias, void* ws, int B,
                             int C, int H, int W, int relu, void* stream);

/* No counterpart in the reference: added for callers that cannot allocate device
 * memory (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  Same operation and
 * results as nconv_b200_relu_bias_bwd, without the `ws` argument: the per-channel fp64 sums go to a static device array
 * inside the library (2 doubles per channel, NCONV_B200_NOALLOC_MAX_C channels = 16 KB per GPU, held until the process
 * exits).  C > NCONV_B200_NOALLOC_MAX_C returns NCONV_B200_ERR_STATIC_SCRATCH.  The call makes no cudaMalloc / cudaFree
 * and calls no allocating library routine; it enqueues a memset of the array, the kernel and (gbias != NULL) the
 * read-out on `stream`.  Every call on a GPU shares the one array: calls that overlap on the GPU (different streams with
 * no dependency between them) race; calls ordered on one stream, or by events, do not. */
#define NCONV_B200_NOALLOC_MAX_C 1024
int nconv_b200_relu_bias_bwd_noalloc(const float* gy, const float* y, float* g, float* gbias, int B, int C, int H,
                                     int W, int relu, void* stream);

/* weight gradient gw[CO][CI][K][K] = sum_{b,oy,ox} g[b,co,oy,ox]*x[b,ci,oy*S-P+u,ox*S-P+v] of
 * Conv2d (K,S,P) in {(1,1,0),(1,2,0),(3,1,1),(3,2,1)}; (4,2,1) with g and x swapped is
 * ConvTranspose2d(4,2,1)'s [CI][CO][4][4] gradient.  ws: split-K partials. */
size_t nconv_b200_conv_wgrad_ws_bytes(int CI, int CO, int K);
int nconv_b200_conv_wgrad(const float* g, const float* x, float* gw, void* ws, size_t ws_bytes, int B,
                          int CI, int CO, int H, int W, int K, int S, int P, void* stream);

/* data gradient of a stride-2 Conv2d (transposed=0, w [CO][CI][K][K]) or of ConvTranspose2d(4,2,1)
 * (transposed=1, w [CI][CO][4][4]); gx: [B,CI,H,W]. */
size_t nconv_b200_conv_dgrad_ws_bytes(int CI, int CO, int K);
int nconv_b200_conv_dgrad_gather(const float* g, const float* w, float* gx, void* ws, int B, int CI,
                                 int CO, int H, int W, int K, int S, int P, int transposed,
                                 void* stream);

/* wf[CI][CO][K][K] = flipped, transposed w[CO][CI][K][K]: nconv_b200_conv2d_fwd(g, wf) is the data
 * gradient of a stride-1 "same" convolution. */
int nconv_b200_wflip(const float* w, float* wf, int CO, int CI, int K, void* stream);

/* F.interpolate(mode='bilinear', align_corners=True) to (HO,WO) — step2.py:252,271 */
int nconv_b200_bilinear_ac_fwd(const float* x, float* y, int planes, int H, int W, int HO, int WO,
                               void* stream);

/* zero rows [0,top) and [H-bottom,H) and cols [0,left) — step2.py:122-124 */
int nconv_b200_zero_border(float* x, int planes, int H, int W, int top, int bottom, int left,
                           void* stream);

/* BatchNorm (eval) -> per-channel scale/shift, optionally absorbing a conv bias:
 * scale = gamma/sqrt(var+eps); shift = beta + (bias-mean)*scale. */
int nconv_b200_bn_fold(const float* gamma, const float* beta, const float* mean, const float* var,
                       const float* conv_bias, float* scale, float* shift, int C, float eps,
                       void* stream);

/* ---- tensor-core path of the RGB-guided stack (step2.py ConvBlock :290-297, Basic2d :178-195) --
 * Activations in "P8": [B][ceil(C/8)][hi|lo][H+2][W+2][8] bf16 with a one-pixel zero ring (see
 * csrc/p8.cuh); buffers are nconv_b200_p8_bytes() long (includes tail slack that tile loads may
 * over-read).  pack/unpack convert from/to contiguous fp32 NCHW (pack also writes the zero ring);
 * kernels write interior pixels only, zero_ring re-establishes the ring of a reused buffer. */
size_t nconv_b200_p8_bytes(int B, int C, int H, int W);
int nconv_b200_p8_pack(const float* x, void* p8, int B, int C, int H, int W, void* stream);
int nconv_b200_p8_unpack(const void* p8, float* x, int B, int C, int H, int W, void* stream);
/* unpack into channels [c0, c0+C) of a contiguous fp32 [B, Ctot, H, W] tensor (the data gradient of a wide layer is computed
 * in blocks of 64 input channels: no concatenation pass afterwards) */
int nconv_b200_p8_unpack_into(const void* p8, float* x, int B, int C, int H, int W, int Ctot, int c0, void* stream);
int nconv_b200_p8_zero_ring(void* p8, int B, int C, int H, int W, void* stream);
/* pack from a channel sub-range of x [B][Ctot][..][..] (channels c0 .. c0+C), optionally with the 2x2 spatial phases of a
 * [2H][2W] source moved to channels (phases != 0: the P8 tensor has 4*C channels of H x W, channel (py*2+px)*C + c at (i, j)
 * = x[c0+c][2i+py][2j+px]; C % 8 == 0).  With the two weight helpers below this turns the backward of Basic2dTrans'
 * ConvTranspose2d(4, 2, 1) (step2.py:197-214 under train_step2.py:62-64) into 3x3 / stride-1 problems for the tcgen05
 * kernels: data gradient = conv3x3(phases(gy), w3), weight gradient = wgrad3x3(phases(gy), x) gathered back. */
int nconv_b200_p8_pack_ex(const float* x, void* p8, int B, int C, int H, int W, int Ctot, int c0, int phases, void* stream);
/* The same device for the backward of a 3x3 / stride-2 / pad-1 Conv2d (RGBEncoder, step2.py:134-154): x[2i+u-1] is phase
 * (u odd ? 0 : 1) of x at row i + {-1, 0}, so  weight gradient = wgrad3x3(phases(x), gy) gathered back (convS2_phase_wgrad),
 * data gradient, phase by phase, = conv3x3(gy, w3d) with w3d from convS2_phase_weights, stored interleaved (y_phase_c).
 *   w [CO][CI][3][3] -> w3d [4*CI][CO][3][3];   gw3 [CO][4*CI][3][3] -> gw [CO][CI][3][3] */
int nconv_b200_convS2_phase_weights(const float* w, float* w3d, int CO, int CI, void* stream);
int nconv_b200_convS2_phase_wgrad(const float* gw3, float* gw, int CO, int CI, void* stream);
/* w [CI][CO][4][4] (ConvTranspose2d layout) -> w3 [CI][4*CO][3][3] */
int nconv_b200_convT_phase_weights(const float* w, float* w3, int CI, int CO, void* stream);
/* gw3 [CIF][4*CO][3][3] -> gw [lead + CIF][CO][4][4], rows lead .. lead+CIF (the first `lead` rows are left alone) */
int nconv_b200_convT_phase_wgrad(const float* gw3, float* gw, int CIF, int CO, int lead, void* stream);
/* Backward of ReLU(conv + bias) (step2.py:290-297 under train_step2.py:62-64) written straight into P8 for the tcgen05
 * data / weight gradient kernels: p8 = P8(gy * [y > 0]) (y null: no mask), gbias[c] = per-channel sum of the masked
 * gradient (null: not wanted).  ws: C doubles of scratch when gbias is given.  gy, y: fp32 NCHW [B,C,H,W]. */
int nconv_b200_relu_bias_bwd_p8(const float* gy, const float* y, void* p8, float* gbias, void* ws, int B, int C, int H,
                                int W, void* stream);

/* P8Q: the four spatial phases of an activation, for stride-2 consumers (csrc/p8.cuh). */
size_t nconv_b200_p8q_bytes(int B, int C, int H, int W);
int nconv_b200_p8q_zero_ring(void* p8q, int B, int C, int H, int W, void* stream);
int nconv_b200_p8q_pack(const float* x, void* p8q, int B, int C, int H, int W, void* stream);

/* Tensor-core convolution modes */
#define NCONV_B200_TC_S1 0 /* 3x3 stride 1 pad 1: ConvBlock step2.py:290-297, Basic2d :178-195       */
#define NCONV_B200_TC_S2 1 /* 3x3 stride 2 pad 1 (+1x1 stride 2 residual): RGBEncoder :134-154       */
#define NCONV_B200_TC_T 2  /* ConvTranspose2d 4x4 stride 2 pad 1: Basic2dTrans :197-214             */

/* weights -> per-K-stage packed split-bf16 image consumed by conv_tc.
 *   S1: w [CO][CA+CB][3][3];  S2: w [CO][CA+CB][3][3] and optionally w1 [CO][CA+CB][1][1];
 *   T : w [CA+CB][CO][4][4] (CA == 1: the lone depth channel of cat[depth, features]). */
size_t nconv_b200_tc_wpack_bytes(int mode, int CO, int CA, int CB, int has_1x1);
int nconv_b200_tc_pack_weights(int mode, const float* w, const float* w1, void* wpack, int CO, int CA,
                               int CB, void* stream);

/* y = act(scale[o]*conv(cat[a,b]) + shift[o]) (+ conv1x1_s2(cat[a,b]) in mode S2 with has_1x1),
 * tcgen05 bf16x3 (fp32-class accuracy).  CA, CB multiples of 16 (mode T: CA may be 1); CO in
 * {32, 64}.  a, b: P8 (S1, T) or P8Q (S2) of size H x W; y: P8 of the output size (S1: HxW,
 * S2: H/2 x W/2, T: 2H x 2W); yq: optional P8Q copy of y for a stride-2 consumer. */
typedef struct {
  const void* a;
  const void* b;        /* concatenated after a, or NULL                        */
  const void* wpack;
  const float* scale;   /* [CO] or NULL (=1)                                   */
  const float* shift;   /* [CO] or NULL (=0)                                   */
  void* y;
  void* yq;             /* or NULL                                             */
  int B, CA, CB, CO, H, W, relu, mode, has_1x1;
  float* y_f32;         /* mode S1 only, or NULL: write fp32 NCHW instead of y, into channels [y_c0, y_c0+CO) of a
                           contiguous [B][y_ctot][H][W] tensor (the data gradient of train_step2.py:62-64's backward);
                           NCONV_B200_ERR_UNSUPPORTED for CO = 32 with CA+CB >= 64 (that shape's kernel writes P8 only) */
  int y_ctot, y_c0;
  int y_phase_c;        /* with y_f32, > 0: the CO output channels are phase-major, channel (py*2+px)*y_phase_c + c, and are
                           written interleaved into y_f32 = [B][y_phase_c][2H][2W] at [c][2y+py][2x+px] (the data gradient of a
                           3x3 / stride-2 convolution computed as a stride-1 convolution of its output gradient); y_c0 = first
                           phase-major channel of this launch, y_phase_c % 32 == 0, y_c0 % 32 == 0; y_ctot unused */
} nconv_b200_tc_desc;
int nconv_b200_conv_tc(const nconv_b200_tc_desc* d, void* stream);

/* HBM-bound edges of the tensor-core stack (CUDA cores, converting fp32 NCHW <-> P8 on the fly).
 * conv_small_p8: y = act(scale*conv3x3(x)+shift) (+ conv1x1(x) when w1 != NULL) for x fp32
 *   [B,CI,H,W] with (CI,CO) in {(3,32),(1,32),(1,64)} -> P8 y (+ optional P8Q copy yq):
 *   rgb_encoder0 (step2.py:134-154) and depth_conv (:221).
 * head_p8: y = residual + conv3x3_{C->1}(f) (Conv3x3 :156, :255-257, :274-276) for f P8
 *   [B,C in {32,64},H,W] -> fp32 y [B,1,H,W] and, optionally, y8: the same plane as a
 *   lone-channel P8 (input of the next level's ConvTranspose). */
int nconv_b200_conv_small_p8(const float* x, const float* w, const float* w1, const float* scale,
                             const float* shift, void* y, void* yq, int B, int CI, int CO, int H, int W,
                             int relu, void* stream);
int nconv_b200_head_p8(const void* f, const float* w, const float* residual, float* y, void* y8, int B,
                       int C, int H, int W, void* stream);

/* ---- callers either side of the training path ------------------------------
 * One scale of utils.py:63-71 (calculate_loss_multi_resolution) over utils.py:138-151
 * (calculate_loss, use_gradient_loss = False):
 *   up = F.interpolate(d, size=(H,W), mode='bilinear', align_corners=False)        d: [N,1,h,w]
 *   loss (+)= weight * mean_{N,H,W}( (up.masked_fill(gt == 0, 0) - gt)^2 )         gt: [N,1,H,W]
 * h == H, w == W is the plain calculate_loss. ws: masked_mse_ws_bytes() of scratch for either
 * call; accumulate != 0 adds to *loss (the mean over scales is weight = 1/nscales).
 * bwd: gd[N,1,h,w] = gloss[0] * dloss/dd  (overwrites gd; deterministic gather, no atomics; for
 * h != H the masked residual at full resolution passes through ws). */
size_t nconv_b200_masked_mse_ws_bytes(int N, int h, int w, int H, int W);
int nconv_b200_masked_mse_bilinear_fwd(const float* d, const float* gt, float* loss, void* ws, size_t ws_bytes,
                                       int N, int h, int w, int H, int W, float weight, int accumulate,
                                       void* stream);
int nconv_b200_masked_mse_bilinear_bwd(const float* d, const float* gt, const float* gloss, float* gd, void* ws,
                                       size_t ws_bytes, int N, int h, int w, int H, int W, float weight,
                                       void* stream);

/* The same with use_gradient_loss = True (utils.py:142-147 over :95-136, the setting of train_step1.py:20):
 *   rec = up.masked_fill(gt == 0, 0)
 *   loss (+)= weight * (0.8*sqrt(mse(rec, gt)) + 0.2*(mean|sobel_x(gt-rec)| + mean|sobel_y(gt-rec)|))
 * (3x3 cross-correlations with zero padding 1).  stats: 3 doubles on the device written by fwd (the three
 * sums) and read by bwd.  ws: sobel_loss_ws_bytes() of scratch for either call.  N <= 65535. */
size_t nconv_b200_sobel_loss_ws_bytes(int N, int h, int w, int H, int W);
int nconv_b200_sobel_loss_fwd(const float* d, const float* gt, float* loss, double* stats, void* ws,
                              size_t ws_bytes, int N, int h, int w, int H, int W, float weight, int accumulate,
                              void* stream);
int nconv_b200_sobel_loss_bwd(const float* d, const float* gt, const float* gloss, const double* stats,
                              float* gd, void* ws, size_t ws_bytes, int N, int h, int w, int H, int W,
                              float weight, void* stream);

/* torch.optim.AdamW (utils.py:53-55 get_optimizer; amsgrad off) on flat fp32 buffers of n elements:
 *   p *= 1 - lr*wd;  m = b1*m + (1-b1)*g;  v = b2*v + (1-b2)*g*g;
 *   p -= lr/(1-b1^step) * m / (sqrt(v)/sqrt(1-b2^step) + eps)                      step >= 1 */
int nconv_b200_adamw_step(float* p, const float* g, float* m, float* v, long long n, float lr, float beta1,
                          float beta2, float eps, float weight_decay, int step, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* NCONV_B200_H_ */
