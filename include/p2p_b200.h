/* p2p_b200.h -- C ABI of libp2p_b200.so: the sm_100a kernels behind the pix2pix G/D training
 * step (define_G / define_D / GANLoss / Adam of the reference).
 *
 * The reference has no FFI of its own: its hot path is `torch.nn` modules dispatching to
 * ATen/cuDNN (reference networks.py, train.py:294-390).  Each entry point below therefore
 * cites the reference *call site* whose ATen/cuDNN work it replaces.  The Python host side
 * (`p2p-pytorch_b200/*.py`) binds these with ctypes; INTEGRATION.md shows the binding.
 *
 * Conventions
 *   - plain pointers are DEVICE pointers unless the name ends in _host; no torch types here;
 *   - nothing here allocates or frees device memory, everything is enqueued on `stream`
 *     (a cudaStream_t passed as void*) and returns without synchronising;
 *   - return value: 0 = enqueued; <0 = rejected argument (see p2p_last_error());
 *     >0 = cudaError_t / CUresult raised while enqueueing;
 *   - activation tensors are NHWC with the channel count padded to a multiple of 16
 *     (pad channels hold zeros); `dtype` selects their element type: P2P_F32 or P2P_BF16;
 *     parameters, gradients of parameters, statistics and losses are always fp32/fp64.
 */
#ifndef P2P_B200_H_
#define P2P_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define P2P_F32 0
#define P2P_BF16 1

/* activations understood by the fused kernels */
#define P2P_ACT_NONE 0
#define P2P_ACT_RELU 1
#define P2P_ACT_PRELU 2  /* one shared learnable slope (reference networks.py:452) */
#define P2P_ACT_LRELU 3  /* fixed slope 0.2 (networks.py:453, :766) */
#define P2P_ACT_TANH 4

/* implementation selector for the convolution entry points */
#define P2P_IMPL_AUTO 0  /* tcgen05 path when the shape is supported and dtype is bf16, else SIMT */
#define P2P_IMPL_SIMT 1  /* fp32-accumulate CUDA-core kernels: the fp32 verification mode */
#define P2P_IMPL_TC 2    /* tcgen05 / TMEM / TMA implicit GEMM; error if the shape is unsupported */

/* One convolution layer as the kernels see it.  Input and output are NHWC tensors *as stored*:
 * reflection / nearest-upsample halos are materialised by the producer (p2p_bn_act_store), so
 * G layers have pad = 0 and Hi = H + 2*(k/2); D layers read the unpadded map with pad = 2 and
 * out-of-range taps contribute zero (TMA zero fill).
 *     out[n,oy,ox,co] = bias[co] + sum_{ky,kx,ci} in[n, oy*stride+ky-pad, ox*stride+kx-pad, ci] * w[co,ky,kx,ci]
 */
typedef struct p2p_conv {
  int32_t N;
  int32_t Hi, Wi, Ci; /* stored input dims, Ci % 16 == 0 */
  int32_t Ho, Wo, Co; /* stored output dims, Co % 16 == 0 */
  int32_t kh, kw, stride, pad;
  int32_t co_real; /* real output channels (<= Co); channels >= co_real are written as 0 */
  int32_t ci_real; /* real input channels (<= Ci) */
  int32_t act;     /* epilogue activation: P2P_ACT_NONE or P2P_ACT_LRELU */
} p2p_conv;

int p2p_version(void);
const char* p2p_last_error(void);
/* 1 if the running device can execute the tcgen05 kernels (compute capability 10.x). */
int p2p_device_supports_tc(void);
/* number of kernels this library has enqueued in this process so far (bench.py's gpu_launches). */
long long p2p_launch_count(void);
/* Diagnostics: between p2p_trace_begin() and p2p_trace_end() every convolution entry point appends one line naming
 * the kernel instantiation (and its block geometry) that served the call; p2p_trace_end copies the text into `buf`
 * (at most cap-1 bytes, NUL-terminated) and returns the size needed.  Tests use it to show that the shapes they
 * check run on the same instantiations as the benchmark.  No reference counterpart. */
void p2p_trace_begin(void);
int p2p_trace_end(char* buf, int cap);

/* ---- convolutions: replace nn.Conv2d fwd/bwd (reference networks.py:400,404 ConvLayer;
 *      :416,422 UpsampleConvLayer; :766-784 NLayerDiscriminator) ------------------------- */

/* w_packed: [Co][kh][kw][Ci] in `dtype`.  stats (optional, may be NULL): double[2*Co], per-channel
 * sum and sum of squares of the pre-activation outputs are ADDED to it (train-mode BatchNorm
 * statistics, networks.py:461-490). */
int p2p_conv_fwd(const p2p_conv* c, const void* in, const void* w_packed, const float* bias, void* out,
                 double* stats, int dtype, int impl, void* stream);
/* din[n,y,x,ci] = sum dout[n,oy,ox,co] * w[co,ky,kx,ci] over the taps that touch (y,x).
 * wt_packed: [Ci][kh][kw][Co] in `dtype` (p2p_pack_weights writes both packings). */
int p2p_conv_dgrad(const p2p_conv* c, const void* dout, const void* wt_packed, void* din, int dtype, int impl,
                   void* stream);
/* dw_acc[ky][kx][ci][co] (fp32, Ci x Co as stored) += sum_pixels in[..ci] * dout[..co]  (split-K atomics;
 * the caller zeroes dw_acc once per optimiser step). */
int p2p_conv_wgrad(const p2p_conv* c, const void* in, const void* dout, float* dw_acc, int dtype, int impl,
                   void* stream);
/* PyTorch weight [co_real][ci_real][kh][kw] fp32, optionally divided by *sigma (spectral norm,
 * networks.py:548), -> w_packed [Co][kh][kw][Ci] and wt_packed [Ci][kh][kw][Co] (either may be NULL). */
int p2p_pack_weights(const float* w, const float* sigma, void* w_packed, void* wt_packed, int co_real, int ci_real,
                     int Co, int Ci, int kh, int kw, int dtype, void* stream);
/* grad[co][ci][kh][kw] += scale * dw_acc[ky][kx][ci][co]   (scale = 1/sigma for spectral-norm layers).
 * dot (optional): *dot += sum dw_acc * w  over real channels, used by p2p_sn_weight_grad. */
int p2p_unpack_wgrad(const float* dw_acc, float* grad, const float* w, const float* sigma, float* dot, int co_real,
                     int ci_real, int Co, int Ci, int kh, int kw, void* stream);
/* spectral-norm correction: grad_wbar[i,k] -= (*dot / sigma^2) * u[i] * v[k]  (reference networks.py:547-548;
 * u, v are read when this runs -- quirk Q1 of SURVEY.md section 8a). */
int p2p_sn_weight_grad(float* grad_wbar, const float* dot, const float* sigma, const float* u, const float* v, int rows,
                       int cols, void* stream);

/* Batched forms of the three calls above: one launch for all layers of a network (the per-layer launches are
 * 5-9 us each and there are ~140 of them per training iteration).  `jobs` is a HOST array of n descriptors with
 * the argument meaning of the single-layer call; it is copied into the launch parameters.  n <= P2P_MULTI_MAX
 * per call.  They have no counterpart of their own in the reference: each job is the same weight preparation /
 * gradient un-packing the reference's conv layers do inside cuDNN. */
#define P2P_MULTI_MAX 24
typedef struct p2p_pack_job {
  const float* w; const float* sigma; void* w_packed; void* wt_packed;
  int32_t co_real, ci_real, Co, Ci, kh, kw;
} p2p_pack_job;
typedef struct p2p_unpack_job {
  const float* dw_acc; float* grad; const float* w; const float* sigma; float* dot;
  int32_t co_real, ci_real, Co, Ci, kh, kw;
} p2p_unpack_job;
typedef struct p2p_sn_grad_job {
  float* grad_wbar; const float* dot; const float* sigma; const float* u; const float* v;
  int32_t rows, cols;
} p2p_sn_grad_job;
int p2p_pack_weights_multi(int n, const p2p_pack_job* jobs, int dtype, void* stream);
int p2p_unpack_wgrad_multi(int n, const p2p_unpack_job* jobs, void* stream);
int p2p_sn_weight_grad_multi(int n, const p2p_sn_grad_job* jobs, void* stream);
/* One power iteration for `n_layers` weight matrices at once (networks.py:537-547):
 * v = l2n(W^T u); u = l2n(W v); sigma = u . (W v).  Arrays of device pointers live in HOST memory
 * (they are kernel arguments).  tmp: device scratch of at least sum(rows+cols) floats. */
#define P2P_SN_MAX_LAYERS 16
int p2p_spectral_norm(int n_layers, const float* const* w_host, float* const* u_host, float* const* v_host,
                      float* const* sigma_host, const int32_t* rows_host, const int32_t* cols_host, float* tmp,
                      void* stream);

/* Allocation-free variant of p2p_spectral_norm.  It has no counterpart in the reference; it exists for
 * callers that cannot allocate device memory when they run the discriminator: hosts that were given a fixed
 * memory budget and laid out all their buffers in advance, and latency-sensitive loops where a
 * cudaMalloc/cudaFree pair would synchronise the device in the middle of a step.  Same arguments and results
 * as p2p_spectral_norm, minus `tmp`: the scratch is a static array inside the library, held by every GPU that
 * loads the library until the process exits (P2P_SN_STATIC_MAX_FLOATS floats = 128 KiB).
 *   - sum(rows + cols) over the layers must not exceed P2P_SN_STATIC_MAX_FLOATS (define_D needs 24,192);
 *     larger requests are rejected (<0, see p2p_last_error()) -- use p2p_spectral_norm with your own buffer;
 *   - the first call on each device looks up the array's device address; make that first call before
 *     starting any stream capture on that device.  No call allocates or frees memory;
 *   - all calls share the one array, and it keeps whatever the previous call left in it; each call clears
 *     the part it accumulates into (a memset enqueued on `stream`) before its first kernel.  Two calls
 *     whose kernels can overlap on the GPU -- different streams, or different host threads -- race on it:
 *     serialise them. */
#define P2P_SN_STATIC_MAX_FLOATS 32768
int p2p_spectral_norm_static(int n_layers, const float* const* w_host, float* const* u_host, float* const* v_host,
                             float* const* sigma_host, const int32_t* rows_host, const int32_t* cols_host,
                             void* stream);

/* ---- generator glue: pixel-unshuffle / upsample / reflection pad / BatchNorm / activations --- */

/* x NCHW fp32 [B,3,H,W] -> conv1 input [B,H+8,W+8,16]: PixelUnshuffle(2) + nearest x2 + ReflectionPad(4)
 * (networks.py:173-187, :457, :494-496, :399). */
int p2p_g_input_prep(const float* x, void* dst, int B, int H, int W, int dtype, void* stream);
/* per-channel sum / sum of squares of a stored map, ADDED to stats[2*C] (used after SIMT convs). */
int p2p_channel_stats(const void* r, double* stats, int64_t pixels, int C, int dtype, void* stream);
/* BatchNorm2d training-mode bookkeeping (networks.py:461 etc.; eps 1e-5, momentum 0.1):
 * from stats -> mean/invstd (saved for backward), scale = gamma*invstd, shift = beta - mean*scale,
 * running_mean/var update (unbiased var), num_batches_tracked += 1.  `sb` = float[4*C]:
 * scale | shift | mean | invstd.  eval mode (training == 0) uses the running buffers instead. */
int p2p_bn_finalize(const double* stats, int64_t count, const float* gamma, const float* beta, float* running_mean,
                    float* running_var, int64_t* num_batches_tracked, float* sb, int c_real, int C, int training,
                    void* stream);
/* y = act(scale*r + shift [+ skip]) written into the NEXT layer's input buffer
 * dst[B, H*up + 2*pad, W*up + 2*pad, C] including its halo (reflection of the upsampled image),
 * or, when out_nchw != NULL, into an NCHW fp32 tensor [B,c_real,H,W] (generator output, networks.py:520).
 * skip (optional) is a padded buffer [B, H + 2*skip_pad, W + 2*skip_pad, C] whose interior is read
 * (ResidualBlock, networks.py:442).  sb == NULL means scale 1, shift 0. */
int p2p_bn_act_store(const void* r, const float* sb, const void* skip, int skip_pad, const float* prelu_a, void* dst,
                     float* out_nchw, int B, int H, int W, int C, int c_real, int act, int up, int pad, int dtype,
                     void* stream);
/* p2p_bn_finalize followed by p2p_bn_act_store (NHWC destination) as ONE launch: same arguments, same results bit for
 * bit; `sb` (4*C floats) is still written for the backward kernels.  The bookkeeping kernel is tiny but sits on the
 * generator's forward chain once per BatchNorm (networks.py:461-490, :433-436). */
int p2p_bn_finalize_act_store(const double* stats, int64_t count, const float* gamma, const float* beta,
                              float* running_mean, float* running_var, int64_t* num_batches_tracked, float* sb,
                              int training, const void* r, const void* skip, int skip_pad, const float* prelu_a, void* dst,
                              int B, int H, int W, int C, int c_real, int act, int up, int pad, int dtype, void* stream);
/* Backward of p2p_bn_act_store + BatchNorm, pass 1.  g = fold(dpad) + g1 + g2 (each optional):
 * dpad is the gradient w.r.t. dst (same padded/upsampled geometry), g1/g2 are compact [B,H,W,C] maps,
 * g_nchw an NCHW fp32 gradient (generator output).  dz = g * act'(z), z = scale*r + shift [+ skip];
 * writes dz [B,H,W,C]; ADDS sum(dz), sum(dz*xhat) to red[2*C] and the PReLU slope gradient to *dprelu. */
int p2p_bn_act_bwd_reduce(const void* dpad, const void* g1, const void* g2, const float* g_nchw, const void* r,
                          const float* sb, const void* skip, int skip_pad, const float* prelu_a, void* dz,
                          double* red, double* dprelu, int B, int H, int W, int C, int c_real, int act, int up,
                          int pad, int dtype, void* stream);
/* pass 2: dr = scale * (dz - mean(dz) - xhat*mean(dz*xhat)); also dgamma += sum(dz*xhat), dbeta += sum(dz),
 * dbias += sum(dr) (the conv bias in front of the BatchNorm).  sb == NULL: dr = dz (no BatchNorm). */
int p2p_bn_bwd_apply(const void* dz, const void* r, const float* sb, const double* red, void* dr, float* dgamma,
                     float* dbeta, float* dbias, int64_t pixels, int C, int c_real, int dtype, void* stream);
int p2p_prelu_grad_finish(const double* dprelu, float* grad, void* stream);

/* ---- discriminator glue (networks.py:732-755, train.py:308,315,336) ------------------------ */

/* cat(a,b) (each NCHW fp32 [B,ca,H,W] / [B,cb,H,W], b may be NULL) -> NHWC fp32 [B,H,W,8] (pad channels 0). */
int p2p_d_input_prep(const float* a, int ca, const float* b, int cb, float* dst, int B, int H, int W, void* stream);
/* AvgPool2d(3, stride 2, padding 1, count_include_pad=False) on NHWC fp32. */
int p2p_avgpool3s2_fwd(const float* in, float* out, int B, int H, int W, int C, void* stream);
/* din[B,H,W,C] += backward of the pool applied to dout[B,Ho,Wo,C]. */
int p2p_avgpool3s2_bwd_add(const float* dout, float* din, int B, int H, int W, int C, void* stream);
/* [P,Cs] fp32 -> [P,Cd] dtype (Cd >= Cs, extra channels zero), and the reverse with accumulate. */
int p2p_cast_pad_channels(const float* in, void* out, int64_t pixels, int Cs, int Cd, int dtype, void* stream);
int p2p_uncast_add(const void* in, float* out, int64_t pixels, int Cs, int Cd, int dtype, void* stream);
/* NHWC fp32 [B,H,W,8] gradient -> NCHW fp32 [B,cb,H,W], channels ca..ca+cb-1 (gradient of fake_b). */
int p2p_d_input_grad(const float* dx, float* db, int ca, int cb, int B, int H, int W, void* stream);
/* dz = (g_a + g_b) * (out > 0 ? 1 : 0.2) for LeakyReLU outputs (act == P2P_ACT_LRELU) or the plain sum;
 * g_a, g_b optional compact maps of `dtype`; also ADDS per-channel sum(dz) to dbias (conv bias gradient). */
int p2p_act_bwd(const void* out, const void* g_a, const void* g_b, void* dz, float* dbias, int64_t pixels, int C,
                int c_real, int act, int dtype, void* stream);
/* Second half of a stride-1 convolution with ONE real output channel (networks.py:783, 512 -> 1) computed as a
 * 1x1 convolution with one output column per filter tap (p2p_conv_fwd on the [taps x Ci] filter row, kh = kw = 1,
 * 16 stored output channels) followed by this gather:
 *   out[n,oy,ox,0] = bias[0] + sum_{ky,kx} t[n, oy+ky-pad, ox+kx-pad, ky*kw+kx]   (zero outside; optional LeakyReLU)
 * t: [N,Hi,Wi,16], out: [N,Ho,Wo,16] with channels 1..15 written as zeros.  The input map is then read once
 * instead of once per tap (46 -> 12 us for the 35x35 map at batch 16). */
int p2p_tap_gather_sum(const void* t, const float* bias, void* out, int N, int Hi, int Wi, int Ho, int Wo, int kh,
                       int kw, int pad, int act, int dtype, void* stream);
/* The data-gradient counterpart: a[n,iy,ix,ky*kw+kx] = dy[n, iy+pad-ky, ix+pad-kx, 0] (zero outside; dy and a have
 * 16 stored channels).  p2p_conv_fwd of `a` with the [Ci x taps] filter matrix (kh = kw = 1, 16 input channels) is
 * the gradient w.r.t. the input of the one-output-channel convolution. */
int p2p_tap_scatter(const void* dy, void* a, int N, int Hi, int Wi, int Ho, int Wo, int kh, int kw, int pad, int dtype,
                    void* stream);
/* Wide filters with very few output channels (ExpandNetwork's last layer: 9x9, 32 -> 3, networks.py:487): the layer
 * runs as a (kh x 1) convolution with kw*co_real "fat" output channels -- weights W'[(kx*co_real+co)][ky][0][ci] =
 * w[co][ci][ky][kx], p2p_conv_fwd with kw = 1 -- followed by the horizontal gather
 *   out[n,y,x,co] = bias[co] + sum_kx t[n, y, x+kx, kx*co_real+co]      t: [N,H,Wt=Wo+kw-1,Ct], out: [N,H,Wo,Co];
 * the data-gradient is the adjoint scatter  a[n,y,xa,kx*co_real+co] = dy[n,y,xa-kx,co]  (a: [N,H,Wa=Wo+kw-1,Ca])
 * followed by p2p_conv_dgrad of the fat convolution.  Nine times fewer tensor-core instructions than one
 * 16-column MMA per filter tap. */
int p2p_kx_gather_sum(const void* t, const float* bias, void* out, int N, int H, int Wt, int Wo, int kw, int co_real,
                      int Ct, int Co, int dtype, void* stream);
int p2p_kx_scatter(const void* dy, void* a, int N, int H, int Wo, int Wa, int kw, int co_real, int Cs, int Ca, int dtype,
                   void* stream);

/* ---- compression front end: CompressionNetwork.forward (networks.py:201-236) ---------------
 * Its three convolutions are p2p_conv_fwd calls (5x5 3->64, 3x3 64->64 with BatchNorm statistics, 3x3 stride 2
 * 64->12) glued with p2p_nchw_to_nhwc / p2p_bn_act_store / p2p_bn_finalize; this is the tail:
 * PixelShuffle(2) of res [B,H/2,W/2,16] (12 real channels), F.normalize over the 3 output channels, + x (NCHW fp32),
 * written to y (net_c's output, may be NULL) and/or, quantised to 2^bits - 1 levels as compress() of
 * generate_dataset.py:29-34 does for train.py:297, to `compressed` (may be NULL); both NCHW fp32 [B,3,H,W]. */
int p2p_c_output(const void* res, const float* x, float* y, float* compressed, int B, int H, int W, int bits, int dtype,
                 void* stream);
/* Zero `bytes` bytes at `p` (16-byte aligned) with 16-byte stores: the split-K weight-gradient accumulators handed to
 * p2p_conv_wgrad must start at zero (what autograd's freshly allocated .grad does for the reference's conv backward). */
int p2p_zero_fill(void* p, int64_t bytes, void* stream);
/* 8-bit images -> the fp32 tensors of the iteration: dst = (src / 255 - 0.5) / 0.5, i.e. ToTensor + Normalize(0.5, 0.5)
 * of dataset.py:19-20,30-40 with torchvision's fp32 operation order (bit-identical), and, when `quant` is not NULL,
 * quant = compress(dst * 0.5 + 0.5, bits) of generate_dataset.py:29-34 (the quantiser in front of the generator,
 * train.py:297 without net_c) in the same pass.  n elements, any layout; buffers 16-byte aligned.  Lets a host ship a
 * batch as bytes (4x less host-to-device traffic than the fp32 tensors the reference's DataLoader hands over). */
int p2p_decode_u8(const uint8_t* src, float* dst, float* quant, int64_t n, int bits, void* stream);
/* ---- VGG19 perceptual loss: VGGLoss / Vgg19, networks.py:18-62, used at train.py:377,395.
 * Its thirteen 3x3 pad-1 convolutions + ReLU are p2p_conv_fwd / p2p_conv_dgrad calls with act = P2P_ACT_RELU in the
 * epilogue (frozen weights: no weight gradient), p2p_act_bwd masks the gradient, the L1 taps reuse p2p_l1_fwd / _bwd;
 * these two replace nn.MaxPool2d(2, 2) (torchvision vgg19.features[4, 9, 18, 27]) and its backward on stored maps
 * [N,H,W,C] (H, W even): `in` is the pooling INPUT; the gradient goes to the first maximum of each 2x2 window. */
int p2p_maxpool2_fwd(const void* in, void* out, int N, int H, int W, int C, int dtype, void* stream);
int p2p_maxpool2_bwd(const void* in, const void* dout, void* din, int N, int H, int W, int C, int dtype, void* stream);
/* stored map [B,H,W,C] dtype <-> NCHW fp32 [B,c_real,H,W] (drop-in API edge). */
int p2p_nhwc_to_nchw(const void* in, float* out, int B, int H, int W, int C, int c_real, int dtype, void* stream);
int p2p_nchw_to_nhwc(const float* in, void* out, int B, int H, int W, int C, int c_real, int dtype, void* stream);

/* ---- objective (networks.py:808-850 GANLoss; train.py:344-351 feature matching; :123-126 TV) -- */

/* *loss += weight * mean((p - target)^2) over the c_real channels of a stored map. */
int p2p_mse_const_fwd(const void* p, float target, float weight, float* loss, int64_t pixels, int C, int c_real,
                      int dtype, void* stream);
/* g = gscale * weight * 2 (p - target) / n  (pad channels 0). */
int p2p_mse_const_bwd(const void* p, float target, float weight, const float* gscale, void* g, int64_t pixels, int C,
                      int c_real, int dtype, void* stream);
/* the same on a flat fp32 array of n elements (GANLoss applied to tensors this library did not produce). */
int p2p_mse_flat_fwd(const float* p, float target, float weight, float* loss, int64_t n, void* stream);
/* *loss += weight * mean((a - b)^2) over two flat fp32 arrays: nn.MSELoss(reduction='mean') of train.py:30, used
 * for phase 3's loss_c (train.py:394).  Forward value only: its gradient reaches net_c through torch.round, i.e.
 * it is identically zero (quirk Q3). */
int p2p_mse_pair_fwd(const float* a, const float* b, float weight, float* loss, int64_t n, void* stream);
/* Evaluation metric of the reference (psnr / tensor2img, train.py:32-42, :60-65): sse[n] += sum over image n of
 * (u8(a) - u8(b))^2 with u8(x) = (uint8) clip(x * 255, 0, 255) (float32 multiply, truncation) -- exact integer
 * sums; PSNR = 10 log10(255^2 * per_image / sse).  a, b: [B, per_image] fp32; sse: B zero-initialised uint64. */
int p2p_psnr_sse(const float* a, const float* b, unsigned long long* sse, int B, int64_t per_image, void* stream);
int p2p_mse_flat_bwd(const float* p, float target, float weight, const float* gscale, float* g, int64_t n,
                     void* stream);
/* *loss += weight * mean|f - r| ; g = gscale * weight * sign(f - r) / n. */
int p2p_l1_fwd(const void* f, const void* r, float weight, float* loss, int64_t pixels, int C, int c_real, int dtype,
               void* stream);
int p2p_l1_bwd(const void* f, const void* r, float weight, const float* gscale, void* g, int64_t pixels, int C,
               int c_real, int dtype, void* stream);
/* total variation of an NCHW fp32 image batch: *loss += mean|dx| + mean|dy|; g += gscale * d loss / d x. */
/* The LSGAN / feature-matching sums over several stored maps in ONE launch (the objective touches 15 maps forward and
 * 15 backward per discriminator pass: networks.py:840-850 over 3 scales, train.py:344-351 over 12 maps).  A job is one
 * p2p_mse_const_* (mode 0: against `target`) or p2p_l1_* (mode 1: against map `r`) call: `coef` = the map's weight (the
 * mean over pixels x c_real is applied inside), `g` = where backward writes the map's gradient (forward ignores it),
 * `blocks` is filled in by the library.  `jobs` is a HOST array, n <= P2P_MULTI_MAX. */
typedef struct p2p_loss_job {
  const void* p; const void* r; void* g;
  float target, coef;
  int64_t pixels;
  int32_t C, c_real, mode, blocks;
} p2p_loss_job;
int p2p_map_loss_multi_fwd(int n, const p2p_loss_job* jobs, float* loss, int dtype, void* stream);
int p2p_map_loss_multi_bwd(int n, const p2p_loss_job* jobs, const float* gscale, int dtype, void* stream);
int p2p_tv_fwd(const float* x, float* loss, int B, int C, int H, int W, void* stream);
int p2p_tv_bwd_add(const float* x, const float* gscale, float* g, int B, int C, int H, int W, void* stream);

/* ---- optimiser: torch.optim.Adam as used at train.py:241-242,386,390 (no weight decay, no amsgrad) */
int p2p_adam_step(float* p, const float* g, float* m, float* v, int64_t n, float lr, float beta1, float beta2,
                  float eps, int step, float grad_scale, void* stream);
/* Same update with the step count (>= 1, already incremented) read from DEVICE memory, so that the launch can
 * be replayed from a CUDA graph while the count advances (torch.optim.Adam's capturable=True mode,
 * torch/optim/adam.py).  Bias corrections are computed on the device in double precision. */
int p2p_adam_step_dev(float* p, const float* g, float* m, float* v, int64_t n, float lr, float beta1, float beta2,
                      float eps, const int* step_dev, float grad_scale, void* stream);
/* As p2p_adam_step_dev, with the learning rate and the gradient scale read from DEVICE memory as well
 * (hyper_dev[0] = lr, hyper_dev[1] = grad_scale): the reference steps an LR scheduler every epoch
 * (train.py:441-443, networks.py:104-125), and a launch replayed from a CUDA graph must see the new value. */
int p2p_adam_step_graph(float* p, const float* g, float* m, float* v, int64_t n, float beta1, float beta2, float eps,
                        const int* step_dev, const float* hyper_dev, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* P2P_B200_H_ */
