// Objective and optimiser kernels: LSGAN MSE against a constant label (reference networks.py:808-850),
// feature-matching L1 (train.py:344-351), total variation (train.py:123-126), fused Adam
// (torch.optim.Adam as configured at train.py:241-242).  All are single-pass, vectorised, warp-shuffle
// reduced and HBM-bound: their roofline is bytes moved / measured copy bandwidth.
#include "common.cuh"

namespace p2p {

// ---- sum over a stored map of f(p[, r]) on the real channels ---------------------------------------
// MODE 0: (p - target)^2      MODE 1: |p - r|
template <typename T, int MODE>
__global__ void __launch_bounds__(256) map_loss_fwd_kernel(const T* __restrict__ p, const T* __restrict__ r,
                                                           float target, float coef, float* __restrict__ loss,
                                                           int64_t pixels, int C, int c_real) {
  const int C8 = C >> 3;
  const int64_t total = pixels * C8;
  float acc = 0.f;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c8 = (int)(i % C8);
    if (c8 * 8 >= c_real) continue;
    float a[8], b[8];
    ld8(p + i * 8, a);
    if (MODE == 1) ld8(r + i * 8, b);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      if (c8 * 8 + k < c_real) {
        const float d = MODE == 0 ? a[k] - target : a[k] - b[k];
        acc += MODE == 0 ? d * d : fabsf(d);
      }
    }
  }
  const float tot = block_sum<float>(acc);
  if (threadIdx.x == 0) atomicAdd(loss, tot * coef);
}

template <typename T, int MODE>
__global__ void __launch_bounds__(256) map_loss_bwd_kernel(const T* __restrict__ p, const T* __restrict__ r,
                                                           float target, float coef, const float* __restrict__ gscale,
                                                           T* __restrict__ g, int64_t pixels, int C, int c_real) {
  const int C8 = C >> 3;
  const int64_t total = pixels * C8;
  const float s = coef * (gscale != nullptr ? *gscale : 1.f);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c8 = (int)(i % C8);
    float a[8], b[8], o[8];
    ld8(p + i * 8, a);
    if (MODE == 1) ld8(r + i * 8, b);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      float v = 0.f;
      if (c8 * 8 + k < c_real) {
        if (MODE == 0) v = 2.f * (a[k] - target) * s;
        else { const float d = a[k] - b[k]; v = d > 0.f ? s : (d < 0.f ? -s : 0.f); }
      }
      o[k] = v;
    }
    st8(g + i * 8, o);
  }
}

// ---- the same two kernels over up to P2P_MULTI_MAX maps in ONE launch (blockIdx.y = map): the objective touches 15
// maps forward and 15 backward per pass, 3-10 us each, in series on the generator's critical path -------------------
struct LossTable {
  p2p_loss_job j[P2P_MULTI_MAX];
};
template <typename T>
__global__ void __launch_bounds__(256) map_loss_multi_fwd_kernel(const __grid_constant__ LossTable tb, float* __restrict__ loss) {
  const p2p_loss_job& jb = tb.j[blockIdx.y];
  const int C8 = jb.C >> 3;
  const int64_t total = jb.pixels * C8;
  if ((int64_t)blockIdx.x * blockDim.x >= total) return;
  const T* p = (const T*)jb.p;
  const T* r = (const T*)jb.r;
  const int64_t stride = (int64_t)jb.blocks * blockDim.x;
  if (blockIdx.x >= (unsigned)jb.blocks) return;
  float acc = 0.f;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int c8 = (int)(i % C8);
    if (c8 * 8 >= jb.c_real) continue;
    float a[8], b[8];
    ld8(p + i * 8, a);
    if (jb.mode == 1) ld8(r + i * 8, b);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      if (c8 * 8 + k < jb.c_real) {
        const float d = jb.mode == 0 ? a[k] - jb.target : a[k] - b[k];
        acc += jb.mode == 0 ? d * d : fabsf(d);
      }
    }
  }
  const float tot = block_sum<float>(acc);
  if (threadIdx.x == 0) atomicAdd(loss, tot * jb.coef);
}

template <typename T>
__global__ void __launch_bounds__(256) map_loss_multi_bwd_kernel(const __grid_constant__ LossTable tb,
                                                                 const float* __restrict__ gscale) {
  const p2p_loss_job& jb = tb.j[blockIdx.y];
  if (blockIdx.x >= (unsigned)jb.blocks) return;
  const int C8 = jb.C >> 3;
  const int64_t total = jb.pixels * C8;
  const T* p = (const T*)jb.p;
  const T* r = (const T*)jb.r;
  T* g = (T*)jb.g;
  const float s = jb.coef * (gscale != nullptr ? *gscale : 1.f);
  const int64_t stride = (int64_t)jb.blocks * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int c8 = (int)(i % C8);
    float a[8], b[8], o[8];
    ld8(p + i * 8, a);
    if (jb.mode == 1) ld8(r + i * 8, b);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      float v = 0.f;
      if (c8 * 8 + k < jb.c_real) {
        if (jb.mode == 0) v = 2.f * (a[k] - jb.target) * s;
        else { const float d = a[k] - b[k]; v = d > 0.f ? s : (d < 0.f ? -s : 0.f); }
      }
      o[k] = v;
    }
    st8(g + i * 8, o);
  }
}

// ---- same on a flat fp32 array (drop-in GANLoss on tensors that did not come from this library) -------
__global__ void __launch_bounds__(256) mse_flat_fwd_kernel(const float* __restrict__ p, float target, float coef,
                                                           float* __restrict__ loss, int64_t n) {
  float acc = 0.f;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float d = p[i] - target;
    acc = fmaf(d, d, acc);
  }
  const float tot = block_sum<float>(acc);
  if (threadIdx.x == 0) atomicAdd(loss, tot * coef);
}
__global__ void mse_flat_bwd_kernel(const float* __restrict__ p, float target, float coef,
                                    const float* __restrict__ gscale, float* __restrict__ g, int64_t n) {
  const float s = 2.f * coef * (gscale != nullptr ? *gscale : 1.f);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    g[i] = (p[i] - target) * s;
}

// ---- mean squared error between two flat fp32 arrays (phase 3's loss_c, train.py:394; forward value only) ----
__global__ void __launch_bounds__(256) mse_pair_fwd_kernel(const float* __restrict__ a, const float* __restrict__ b,
                                                           float coef, float* __restrict__ loss, int64_t n) {
  float acc = 0.f;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float d = a[i] - b[i];
    acc = fmaf(d, d, acc);
  }
  const float tot = block_sum<float>(acc);
  if (threadIdx.x == 0) atomicAdd(loss, tot * coef);
}

// ---- evaluation metric (train.py:32-42, :60-65): sum of squared differences of the two images as uint8 --------
// u8(x) = (uint8) clip(x * 255, 0, 255) -- float32 multiply, truncation -- so the sum is exact integer work
__global__ void __launch_bounds__(256) psnr_sse_kernel(const float* __restrict__ a, const float* __restrict__ b,
                                                       unsigned long long* __restrict__ sse, int64_t per_image) {
  const float* pa = a + (size_t)blockIdx.y * per_image;
  const float* pb = b + (size_t)blockIdx.y * per_image;
  unsigned long long acc = 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < per_image; i += (int64_t)gridDim.x * blockDim.x) {
    const int qa = (int)fminf(fmaxf(pa[i] * 255.f, 0.f), 255.f), qb = (int)fminf(fmaxf(pb[i] * 255.f, 0.f), 255.f);
    const int d = qa - qb;
    acc += (unsigned long long)(d * d);
  }
  const unsigned long long tot = block_sum<unsigned long long>(acc);
  if (threadIdx.x == 0 && tot != 0) atomicAdd(sse + blockIdx.y, tot);
}

// ---- total variation on NCHW fp32 -------------------------------------------------------------------
__global__ void __launch_bounds__(256) tv_fwd_kernel(const float* __restrict__ x, float* __restrict__ loss, int64_t planes,
                                                     int H, int W, float cw, float ch) {
  const int64_t total = planes * H * W;
  float acc = 0.f;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int xx = (int)(i % W);
    const int yy = (int)((i / W) % H);
    const float v = x[i];
    if (xx + 1 < W) acc += fabsf(v - x[i + 1]) * cw;
    if (yy + 1 < H) acc += fabsf(v - x[i + W]) * ch;
  }
  const float tot = block_sum<float>(acc);
  if (threadIdx.x == 0) atomicAdd(loss, tot);
}

__device__ __forceinline__ float sgn(float d) { return d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f); }

__global__ void tv_bwd_add_kernel(const float* __restrict__ x, const float* __restrict__ gscale, float* __restrict__ g,
                                  int64_t planes, int H, int W, float cw, float ch) {
  const int64_t total = planes * H * W;
  const float s = gscale != nullptr ? *gscale : 1.f;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int xx = (int)(i % W);
    const int yy = (int)((i / W) % H);
    const float v = x[i];
    float d = 0.f;
    if (xx + 1 < W) d += sgn(v - x[i + 1]) * cw;
    if (xx > 0) d -= sgn(x[i - 1] - v) * cw;
    if (yy + 1 < H) d += sgn(v - x[i + W]) * ch;
    if (yy > 0) d -= sgn(x[i - W] - v) * ch;
    g[i] += d * s;
  }
}

// ---- Adam ------------------------------------------------------------------------------------------
// torch.optim.Adam (no amsgrad, no weight decay): m = b1 m + (1-b1) g; v = b2 v + (1-b2) g^2;
// p -= (lr / (1-b1^t)) * m / (sqrt(v)/sqrt(1-b2^t) + eps).  28 bytes per parameter (read p,g,m,v; write p,m,v).
__global__ void __launch_bounds__(256) adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                                                   float* __restrict__ v, int64_t n4, int64_t n, float b1, float b2,
                                                   float step_size, float inv_sqrt_bc2, float eps, float gs,
                                                   const int* __restrict__ step_dev,
                                                   const float* __restrict__ hyper_dev) {
  if (hyper_dev != nullptr) {
    // learning rate and gradient scale live in device memory too: a scheduler may change them between replays
    step_size = hyper_dev[0];
    gs = hyper_dev[1];
  }
  if (step_dev != nullptr) {
    // step count lives in device memory (the launch is replayed from a CUDA graph): same double-precision bias
    // corrections as the host path; step_size arrives holding lr
    const int step = *step_dev;
    step_size = (float)((double)step_size / (1.0 - pow((double)b1, (double)step)));
    inv_sqrt_bc2 = (float)(1.0 / sqrt(1.0 - pow((double)b2, (double)step)));
  }
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    float4 P = reinterpret_cast<float4*>(p)[i];
    const float4 G = reinterpret_cast<const float4*>(g)[i];
    float4 M = reinterpret_cast<float4*>(m)[i];
    float4 V = reinterpret_cast<float4*>(v)[i];
#define P2P_ADAM1(f)                                                   \
  {                                                                    \
    const float gg = G.f * gs;                                         \
    M.f = M.f + (1.f - b1) * (gg - M.f);                               \
    V.f = b2 * V.f + (1.f - b2) * gg * gg;                             \
    P.f -= step_size * (M.f / (sqrtf(V.f) * inv_sqrt_bc2 + eps));      \
  }
    P2P_ADAM1(x) P2P_ADAM1(y) P2P_ADAM1(z) P2P_ADAM1(w)
#undef P2P_ADAM1
    reinterpret_cast<float4*>(p)[i] = P;
    reinterpret_cast<float4*>(m)[i] = M;
    reinterpret_cast<float4*>(v)[i] = V;
  }
  // tail (n not a multiple of 4)
  const int64_t t = n4 * 4 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t < n) {
    const float gg = g[t] * gs;
    const float mm = m[t] + (1.f - b1) * (gg - m[t]);
    const float vv = b2 * v[t] + (1.f - b2) * gg * gg;
    m[t] = mm;
    v[t] = vv;
    p[t] -= step_size * (mm / (sqrtf(vv) * inv_sqrt_bc2 + eps));
  }
}

static inline int red_grid(int64_t total) {
  int64_t g = (total + 255) / 256;
  const int64_t cap = (int64_t)kSMs * 8;
  return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace p2p

using namespace p2p;

extern "C" {

int p2p_mse_const_fwd(const void* p, float target, float weight, float* loss, int64_t pixels, int C, int c_real,
                      int dtype, void* stream) {
  P2P_REQUIRE(p && loss && pixels > 0 && C % 8 == 0 && c_real <= C, "bad arguments");
  const float coef = weight / (float)((double)pixels * c_real);
  P2P_DISPATCH(dtype, map_loss_fwd_kernel<T, 0><<<red_grid(pixels * (C / 8)), 256, 0, as_stream(stream)>>>(
                          (const T*)p, nullptr, target, coef, loss, pixels, C, c_real););
  P2P_LAUNCH_CHECK();
}

int p2p_mse_const_bwd(const void* p, float target, float weight, const float* gscale, void* g, int64_t pixels, int C,
                      int c_real, int dtype, void* stream) {
  P2P_REQUIRE(p && g && pixels > 0 && C % 8 == 0 && c_real <= C, "bad arguments");
  const float coef = weight / (float)((double)pixels * c_real);
  P2P_DISPATCH(dtype, map_loss_bwd_kernel<T, 0><<<red_grid(pixels * (C / 8)), 256, 0, as_stream(stream)>>>(
                          (const T*)p, nullptr, target, coef, gscale, (T*)g, pixels, C, c_real););
  P2P_LAUNCH_CHECK();
}

int p2p_mse_flat_fwd(const float* p, float target, float weight, float* loss, int64_t n, void* stream) {
  P2P_REQUIRE(p && loss && n > 0, "bad arguments");
  mse_flat_fwd_kernel<<<red_grid(n), 256, 0, as_stream(stream)>>>(p, target, weight / (float)n, loss, n);
  P2P_LAUNCH_CHECK();
}

int p2p_mse_pair_fwd(const float* a, const float* b, float weight, float* loss, int64_t n, void* stream) {
  P2P_REQUIRE(a && b && loss && n > 0, "bad arguments");
  mse_pair_fwd_kernel<<<red_grid(n), 256, 0, as_stream(stream)>>>(a, b, weight / (float)n, loss, n);
  P2P_LAUNCH_CHECK();
}

int p2p_psnr_sse(const float* a, const float* b, unsigned long long* sse, int B, int64_t per_image, void* stream) {
  P2P_REQUIRE(a && b && sse && B > 0 && B <= 65535 && per_image > 0, "bad arguments");
  int64_t gx = (per_image + 255) / 256;
  if (gx > 64) gx = 64;
  psnr_sse_kernel<<<dim3((int)gx, B), 256, 0, as_stream(stream)>>>(a, b, sse, per_image);
  P2P_LAUNCH_CHECK();
}

int p2p_mse_flat_bwd(const float* p, float target, float weight, const float* gscale, float* g, int64_t n,
                     void* stream) {
  P2P_REQUIRE(p && g && n > 0, "bad arguments");
  mse_flat_bwd_kernel<<<red_grid(n), 256, 0, as_stream(stream)>>>(p, target, weight / (float)n, gscale, g, n);
  P2P_LAUNCH_CHECK();
}

int p2p_l1_fwd(const void* f, const void* r, float weight, float* loss, int64_t pixels, int C, int c_real, int dtype,
               void* stream) {
  P2P_REQUIRE(f && r && loss && pixels > 0 && C % 8 == 0 && c_real <= C, "bad arguments");
  const float coef = weight / (float)((double)pixels * c_real);
  P2P_DISPATCH(dtype, map_loss_fwd_kernel<T, 1><<<red_grid(pixels * (C / 8)), 256, 0, as_stream(stream)>>>(
                          (const T*)f, (const T*)r, 0.f, coef, loss, pixels, C, c_real););
  P2P_LAUNCH_CHECK();
}

int p2p_l1_bwd(const void* f, const void* r, float weight, const float* gscale, void* g, int64_t pixels, int C,
               int c_real, int dtype, void* stream) {
  P2P_REQUIRE(f && r && g && pixels > 0 && C % 8 == 0 && c_real <= C, "bad arguments");
  const float coef = weight / (float)((double)pixels * c_real);
  P2P_DISPATCH(dtype, map_loss_bwd_kernel<T, 1><<<red_grid(pixels * (C / 8)), 256, 0, as_stream(stream)>>>(
                          (const T*)f, (const T*)r, 0.f, coef, gscale, (T*)g, pixels, C, c_real););
  P2P_LAUNCH_CHECK();
}

static int fill_loss_table(int n, const p2p_loss_job* jobs, LossTable* tb, int* max_blocks) {
  P2P_REQUIRE(n >= 1 && n <= P2P_MULTI_MAX && jobs != nullptr, "1..P2P_MULTI_MAX jobs expected");
  *max_blocks = 1;
  for (int i = 0; i < n; ++i) {
    p2p_loss_job j = jobs[i];
    P2P_REQUIRE(j.p && j.pixels > 0 && j.C % 8 == 0 && j.c_real <= j.C && (j.mode == 0 || (j.mode == 1 && j.r)),
                "bad loss job");
    // coef arrives as the job's weight; the mean over the real elements is folded in here, like the single-map calls
    j.coef = j.coef / (float)((double)j.pixels * j.c_real);
    j.blocks = red_grid(j.pixels * (j.C / 8));
    if (j.blocks > *max_blocks) *max_blocks = j.blocks;
    tb->j[i] = j;
  }
  return 0;
}

int p2p_map_loss_multi_fwd(int n, const p2p_loss_job* jobs, float* loss, int dtype, void* stream) {
  LossTable tb{};
  int mb = 1;
  if (int e = fill_loss_table(n, jobs, &tb, &mb)) return e;
  P2P_REQUIRE(loss != nullptr, "null pointer");
  P2P_DISPATCH(dtype, map_loss_multi_fwd_kernel<T><<<dim3(mb, n), 256, 0, as_stream(stream)>>>(tb, loss););
  P2P_LAUNCH_CHECK();
}

int p2p_map_loss_multi_bwd(int n, const p2p_loss_job* jobs, const float* gscale, int dtype, void* stream) {
  LossTable tb{};
  int mb = 1;
  if (int e = fill_loss_table(n, jobs, &tb, &mb)) return e;
  for (int i = 0; i < n; ++i) P2P_REQUIRE(jobs[i].g != nullptr, "null gradient pointer");
  P2P_DISPATCH(dtype, map_loss_multi_bwd_kernel<T><<<dim3(mb, n), 256, 0, as_stream(stream)>>>(tb, gscale););
  P2P_LAUNCH_CHECK();
}

int p2p_tv_fwd(const float* x, float* loss, int B, int C, int H, int W, void* stream) {
  P2P_REQUIRE(x && loss && H > 1 && W > 1, "bad arguments");
  const int64_t planes = (int64_t)B * C;
  const float cw = 1.f / (float)((double)planes * H * (W - 1)), ch = 1.f / (float)((double)planes * (H - 1) * W);
  tv_fwd_kernel<<<red_grid(planes * H * W), 256, 0, as_stream(stream)>>>(x, loss, planes, H, W, cw, ch);
  P2P_LAUNCH_CHECK();
}

int p2p_tv_bwd_add(const float* x, const float* gscale, float* g, int B, int C, int H, int W, void* stream) {
  P2P_REQUIRE(x && g && H > 1 && W > 1, "bad arguments");
  const int64_t planes = (int64_t)B * C;
  const float cw = 1.f / (float)((double)planes * H * (W - 1)), ch = 1.f / (float)((double)planes * (H - 1) * W);
  tv_bwd_add_kernel<<<red_grid(planes * H * W), 256, 0, as_stream(stream)>>>(x, gscale, g, planes, H, W, cw, ch);
  P2P_LAUNCH_CHECK();
}

int p2p_adam_step(float* p, const float* g, float* m, float* v, int64_t n, float lr, float beta1, float beta2,
                  float eps, int step, float grad_scale, void* stream) {
  P2P_REQUIRE(p && g && m && v && n > 0 && step >= 1, "bad arguments");
  P2P_REQUIRE(((uintptr_t)p % 16 == 0) && ((uintptr_t)g % 16 == 0) && ((uintptr_t)m % 16 == 0) &&
                  ((uintptr_t)v % 16 == 0),
              "buffers must be 16-byte aligned");
  const double bc1 = 1.0 - pow((double)beta1, step), bc2 = 1.0 - pow((double)beta2, step);
  const float step_size = (float)((double)lr / bc1);
  const float inv_sqrt_bc2 = (float)(1.0 / sqrt(bc2));
  const int64_t n4 = n / 4;
  int64_t blocks = (n4 + 255) / 256;
  if (blocks < 1) blocks = 1;
  if (blocks > kSMs * 16) blocks = kSMs * 16;
  adam_kernel<<<(int)blocks, 256, 0, as_stream(stream)>>>(p, g, m, v, n4, n, beta1, beta2, step_size, inv_sqrt_bc2, eps,
                                                         grad_scale, nullptr, nullptr);
  P2P_LAUNCH_CHECK();
}

int p2p_adam_step_dev(float* p, const float* g, float* m, float* v, int64_t n, float lr, float beta1, float beta2,
                      float eps, const int* step_dev, float grad_scale, void* stream) {
  P2P_REQUIRE(p && g && m && v && n > 0 && step_dev, "bad arguments");
  P2P_REQUIRE(((uintptr_t)p % 16 == 0) && ((uintptr_t)g % 16 == 0) && ((uintptr_t)m % 16 == 0) &&
                  ((uintptr_t)v % 16 == 0),
              "buffers must be 16-byte aligned");
  const int64_t n4 = n / 4;
  int64_t blocks = (n4 + 255) / 256;
  if (blocks < 1) blocks = 1;
  if (blocks > kSMs * 16) blocks = kSMs * 16;
  adam_kernel<<<(int)blocks, 256, 0, as_stream(stream)>>>(p, g, m, v, n4, n, beta1, beta2, lr, 1.f, eps, grad_scale,
                                                         step_dev, nullptr);
  P2P_LAUNCH_CHECK();
}

int p2p_adam_step_graph(float* p, const float* g, float* m, float* v, int64_t n, float beta1, float beta2, float eps,
                        const int* step_dev, const float* hyper_dev, void* stream) {
  P2P_REQUIRE(p && g && m && v && n > 0 && step_dev && hyper_dev, "bad arguments");
  P2P_REQUIRE(((uintptr_t)p % 16 == 0) && ((uintptr_t)g % 16 == 0) && ((uintptr_t)m % 16 == 0) &&
                  ((uintptr_t)v % 16 == 0),
              "buffers must be 16-byte aligned");
  const int64_t n4 = n / 4;
  int64_t blocks = (n4 + 255) / 256;
  if (blocks < 1) blocks = 1;
  if (blocks > kSMs * 16) blocks = kSMs * 16;
  adam_kernel<<<(int)blocks, 256, 0, as_stream(stream)>>>(p, g, m, v, n4, n, beta1, beta2, 0.f, 1.f, eps, 1.f, step_dev,
                                                         hyper_dev);
  P2P_LAUNCH_CHECK();
}

}  // extern "C"
