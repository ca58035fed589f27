// HBM-bound glue around the convolutions: layout changes at the API edge, reflection / upsample halos,
// train-mode BatchNorm (statistics finalisation, apply, backward), activations and their gradients,
// the discriminator's input pyramid.  All kernels move 8 channels (16-32 bytes) per thread, channel-fastest,
// so every warp access is a run of full 32-byte sectors.
#include <type_traits>

#include "common.cuh"

namespace p2p {


// =================================================================================================
// generator input: PixelUnshuffle(2) -> nearest x2 -> ReflectionPad(4)   (networks.py:173-187,457,399)
// =================================================================================================
template <typename T>
__global__ void g_input_prep_kernel(const float* __restrict__ x, T* __restrict__ dst, int B, int H, int W) {
  const int Hp = H + 8, Wp = W + 8;
  const int64_t total = (int64_t)B * Hp * Wp * 2;  // two 8-channel groups
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int g = (int)(i & 1);
    int64_t p = i >> 1;
    const int px = (int)(p % Wp); p /= Wp;
    const int py = (int)(p % Hp);
    const int n = (int)(p / Hp);
    const int uy = reflect_idx(py - 4, H), ux = reflect_idx(px - 4, W);   // position in the upsampled image
    const int iy = uy >> 1, ix = ux >> 1;                                  // position in the unshuffled image
    float v[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const int c = g * 8 + k;                                             // c = 4*c3 + 2*dy + dx
      float val = 0.f;
      if (c < 12) {
        const int c3 = c >> 2, dy = (c >> 1) & 1, dx = c & 1;
        val = x[(((int64_t)n * 3 + c3) * H + 2 * iy + dy) * W + 2 * ix + dx];
      }
      v[k] = val;
    }
    st8(dst + i * 8, v);
  }
}

// =================================================================================================
// per-channel statistics of a stored map (after SIMT convs; the tcgen05 epilogue produces them itself)
// =================================================================================================
template <typename T>
__global__ void __launch_bounds__(256) channel_stats_kernel(const T* __restrict__ r, double* __restrict__ stats,
                                                            int64_t pixels, int C) {
  const int C8 = C >> 3;
  const int c8 = threadIdx.x % C8, lane_p = threadIdx.x / C8, ppb = 256 / C8;
  float s[8], q[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) s[k] = q[k] = 0.f;
  for (int64_t p = (int64_t)blockIdx.x * ppb + lane_p; p < pixels; p += (int64_t)gridDim.x * ppb) {
    float v[8];
    ld8(r + (size_t)p * C + c8 * 8, v);
#pragma unroll
    for (int k = 0; k < 8; ++k) { s[k] += v[k]; q[k] += v[k] * v[k]; }
  }
  __shared__ float sh[2][8][256];
#pragma unroll
  for (int k = 0; k < 8; ++k) { sh[0][k][threadIdx.x] = s[k]; sh[1][k][threadIdx.x] = q[k]; }
  __syncthreads();
  // thread t < 2*C sums one (which, channel) column over the ppb pixel lanes
  for (int t = threadIdx.x; t < 2 * C; t += 256) {
    const int which = t / C, c = t % C;
    double a = 0.0;
    for (int l = 0; l < ppb; ++l) a += (double)sh[which][c & 7][l * C8 + (c >> 3)];
    atomicAdd(stats + which * C + c, a);
  }
}

// =================================================================================================
// BatchNorm bookkeeping
// =================================================================================================
// one channel of the BatchNorm bookkeeping: batch statistics (training) or running statistics (eval) -> scale / shift /
// mean / 1/std; `update` (one thread per channel, once per layer) also moves the running buffers
__device__ __forceinline__ void bn_channel(const double* __restrict__ stats, int64_t count, const float* __restrict__ gamma,
                                           const float* __restrict__ beta, float* running_mean, float* running_var, int c,
                                           int c_real, int C, int training, bool update, float out[4]) {
  float scale = 0.f, shift = 0.f, mean = 0.f, invstd = 0.f;
  if (c < c_real) {
    if (training) {
      const double m = stats[c] / (double)count;
      double var = stats[C + c] / (double)count - m * m;
      var = var < 0.0 ? 0.0 : var;
      mean = (float)m;
      invstd = (float)(1.0 / sqrt(var + 1e-5));
      if (update && running_mean != nullptr) {
        const double unb = count > 1 ? var * (double)count / (double)(count - 1) : var;
        running_mean[c] = 0.9f * running_mean[c] + 0.1f * mean;
        running_var[c] = 0.9f * running_var[c] + 0.1f * (float)unb;
      }
    } else {
      mean = running_mean[c];
      invstd = rsqrtf(running_var[c] + 1e-5f);
    }
    scale = gamma[c] * invstd;
    shift = beta[c] - mean * scale;
  }
  out[0] = scale; out[1] = shift; out[2] = mean; out[3] = invstd;
}

__global__ void bn_finalize_kernel(const double* __restrict__ stats, int64_t count, const float* __restrict__ gamma,
                                   const float* __restrict__ beta, float* running_mean, float* running_var,
                                   int64_t* nbt, float* __restrict__ sb, int c_real, int C, int training) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c == 0 && training && nbt != nullptr) *nbt += 1;
  if (c >= C) return;
  float o[4];
  bn_channel(stats, count, gamma, beta, running_mean, running_var, c, c_real, C, training, true, o);
  sb[c] = o[0];
  sb[C + c] = o[1];
  sb[2 * C + c] = o[2];
  sb[3 * C + c] = o[3];
}

__device__ __forceinline__ float act_fwd(float z, int act, float a) {
  switch (act) {
    case P2P_ACT_RELU: return z > 0.f ? z : 0.f;
    case P2P_ACT_PRELU: return z > 0.f ? z : a * z;
    case P2P_ACT_LRELU: return z > 0.f ? z : 0.2f * z;
    case P2P_ACT_TANH: return tanhf(z);
    default: return z;
  }
}
__device__ __forceinline__ float act_bwd(float z, int act, float a) {
  switch (act) {
    case P2P_ACT_RELU: return z > 0.f ? 1.f : 0.f;
    case P2P_ACT_PRELU: return z > 0.f ? 1.f : a;
    case P2P_ACT_LRELU: return z > 0.f ? 1.f : 0.2f;
    case P2P_ACT_TANH: { const float t = tanhf(z); return 1.f - t * t; }
    default: return 1.f;
  }
}

// y = act(scale*r + shift [+ skip]) -> padded / upsampled destination (gather form: one thread per
// destination pixel x 8 channels, source found through the reflection and the >> log2(up)).
template <typename T, int ACT>
__global__ void bn_act_store_kernel(const T* __restrict__ r, const float* __restrict__ sb, const T* __restrict__ skip,
                                    int skip_pad, const float* __restrict__ prelu_a, T* __restrict__ dst, int B, int H,
                                    int W, int C, int up, int pad) {
  constexpr int act = ACT;
  const int C8 = C >> 3, Hu = H * up, Wu = W * up, Hd = Hu + 2 * pad, Wd = Wu + 2 * pad;
  const int sh = up == 2 ? 1 : 0;
  const float a = prelu_a != nullptr ? *prelu_a : 0.f;
  const uint32_t total = (uint32_t)B * Hd * Wd * C8;   // < 2^31, checked by the caller
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int c8 = (int)(i % (uint32_t)C8);
    uint32_t p = i / (uint32_t)C8;
    const int dx = (int)(p % (uint32_t)Wd); p /= (uint32_t)Wd;
    const int dy = (int)(p % (uint32_t)Hd);
    const int n = (int)(p / (uint32_t)Hd);
    const int sy = reflect_idx(dy - pad, Hu) >> sh, sx = reflect_idx(dx - pad, Wu) >> sh;
    float v[8];
    ld8(r + (((int64_t)n * H + sy) * W + sx) * C + c8 * 8, v);
    float sk[8];
    if (skip != nullptr)
      ld8(skip + (((int64_t)n * (H + 2 * skip_pad) + sy + skip_pad) * (W + 2 * skip_pad) + sx + skip_pad) * C + c8 * 8, sk);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const int c = c8 * 8 + k;
      float z = sb != nullptr ? fmaf(v[k], sb[c], sb[C + c]) : v[k];
      if (skip != nullptr) z += sk[k];
      v[k] = act_fwd(z, act, a);
    }
    st8(dst + (size_t)i * 8, v);
  }
}

// bn_finalize + bn_act_store in one launch (the tiny bookkeeping kernel sits on the generator's critical path 24 times
// per forward): every block derives scale / shift of all C <= 512 channels from the statistics into shared memory --
// the same arithmetic as bn_finalize_kernel, so the results are bit-identical -- and block 0 also writes `sb` for the
// backward pass and moves the running buffers.
template <typename T, int ACT>
__global__ void __launch_bounds__(256)
    bn_finalize_act_store_kernel(const double* __restrict__ stats, int64_t count, const float* __restrict__ gamma,
                                 const float* __restrict__ beta, float* running_mean, float* running_var, int64_t* nbt,
                                 float* __restrict__ sb, int c_real, int training, const T* __restrict__ r,
                                 const T* __restrict__ skip, int skip_pad, const float* __restrict__ prelu_a,
                                 T* __restrict__ dst, int B, int H, int W, int C, int up, int pad) {
  __shared__ float s_scale[512], s_shift[512];
  const bool first = blockIdx.x == 0;
  if (first && threadIdx.x == 0 && training && nbt != nullptr) *nbt += 1;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float o[4];
    bn_channel(stats, count, gamma, beta, running_mean, running_var, c, c_real, C, training, first, o);
    s_scale[c] = o[0];
    s_shift[c] = o[1];
    if (first) {
      sb[c] = o[0];
      sb[C + c] = o[1];
      sb[2 * C + c] = o[2];
      sb[3 * C + c] = o[3];
    }
  }
  __syncthreads();
  constexpr int act = ACT;
  const int C8 = C >> 3, Hu = H * up, Wu = W * up, Hd = Hu + 2 * pad, Wd = Wu + 2 * pad;
  const int sh = up == 2 ? 1 : 0;
  const float a = prelu_a != nullptr ? *prelu_a : 0.f;
  const uint32_t total = (uint32_t)B * Hd * Wd * C8;   // < 2^31, checked by the caller
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int c8 = (int)(i % (uint32_t)C8);
    uint32_t p = i / (uint32_t)C8;
    const int dx = (int)(p % (uint32_t)Wd); p /= (uint32_t)Wd;
    const int dy = (int)(p % (uint32_t)Hd);
    const int n = (int)(p / (uint32_t)Hd);
    const int sy = reflect_idx(dy - pad, Hu) >> sh, sx = reflect_idx(dx - pad, Wu) >> sh;
    float v[8];
    ld8(r + (((int64_t)n * H + sy) * W + sx) * C + c8 * 8, v);
    float sk[8];
    if (skip != nullptr)
      ld8(skip + (((int64_t)n * (H + 2 * skip_pad) + sy + skip_pad) * (W + 2 * skip_pad) + sx + skip_pad) * C + c8 * 8, sk);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const int c = c8 * 8 + k;
      float z = fmaf(v[k], s_scale[c], s_shift[c]);
      if (skip != nullptr) z += sk[k];
      v[k] = act_fwd(z, act, a);
    }
    st8(dst + (size_t)i * 8, v);
  }
}

// generator output: y = act(scale*r + shift) -> NCHW fp32 [B,c_real,H,W]
template <typename T>
__global__ void bn_act_store_nchw_kernel(const T* __restrict__ r, const float* __restrict__ sb,
                                         const float* __restrict__ prelu_a, float* __restrict__ out, int B, int H, int W,
                                         int C, int c_real, int act) {
  const float a = prelu_a != nullptr ? *prelu_a : 0.f;
  const int64_t total = (int64_t)B * c_real * H * W;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    int64_t p = i;
    const int x = (int)(p % W); p /= W;
    const int y = (int)(p % H); p /= H;
    const int c = (int)(p % c_real);
    const int n = (int)(p / c_real);
    const float v = ldf(r + (((int64_t)n * H + y) * W + x) * C + c);
    const float z = sb != nullptr ? fmaf(v, sb[c], sb[C + c]) : v;
    out[i] = act_fwd(z, act, a);
  }
}

// raw 8-channel vector: lets a thread issue all of its loads before the first conversion / use
template <typename T> struct Raw8;
template <> struct Raw8<float> {
  float4 a, b;
  __device__ __forceinline__ void load(const float* p) {
    a = reinterpret_cast<const float4*>(p)[0];
    b = reinterpret_cast<const float4*>(p)[1];
  }
  __device__ __forceinline__ void zero() { a = make_float4(0.f, 0.f, 0.f, 0.f); b = a; }
  __device__ __forceinline__ void get(float v[8]) const {
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
  }
};
template <> struct Raw8<__nv_bfloat16> {
  uint4 u;
  __device__ __forceinline__ void load(const __nv_bfloat16* p) { u = *reinterpret_cast<const uint4*>(p); }
  __device__ __forceinline__ void zero() { u = make_uint4(0u, 0u, 0u, 0u); }
  __device__ __forceinline__ void get(float v[8]) const {
    const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      v[2 * i] = __uint_as_float(w[i] << 16);
      v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
    }
  }
};

// rows of the padded buffer that mirror onto upsampled row u besides u+pad itself (-1 = none)
__device__ __forceinline__ void fold_extra(int u, int n, int pad, int out[2]) {
  out[0] = (u >= 1 && u <= pad) ? pad - u : -1;                                   // top / left halo
  out[1] = (u <= n - 2 && u >= n - 1 - pad) ? 2 * (n - 1) - u + pad : -1;         // bottom / right halo
}

// raw 4-channel vector
template <typename T> struct Raw4;
template <> struct Raw4<float> {
  float4 a;
  __device__ __forceinline__ void load(const float* p) { a = *reinterpret_cast<const float4*>(p); }
  __device__ __forceinline__ void get(float v[4]) const { v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; }
  static __device__ __forceinline__ void store(float* p, const float v[4]) {
    *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
  }
};
template <> struct Raw4<__nv_bfloat16> {
  uint2 u;
  __device__ __forceinline__ void load(const __nv_bfloat16* p) { u = *reinterpret_cast<const uint2*>(p); }
  __device__ __forceinline__ void get(float v[4]) const {
    v[0] = __uint_as_float(u.x << 16); v[1] = __uint_as_float(u.x & 0xffff0000u);
    v[2] = __uint_as_float(u.y << 16); v[3] = __uint_as_float(u.y & 0xffff0000u);
  }
  static __device__ __forceinline__ void store(__nv_bfloat16* p, const float v[4]) {
    __nv_bfloat162 lo = __floats2bfloat162_rn(v[0], v[1]), hi = __floats2bfloat162_rn(v[2], v[3]);
    *reinterpret_cast<uint2*>(p) = make_uint2(*reinterpret_cast<uint32_t*>(&lo), *reinterpret_cast<uint32_t*>(&hi));
  }
};

// Mirrored copies of source pixel (n, y, x) in the reflection halo of the padded gradient buffer, 4 channels.
template <typename T>
__device__ __forceinline__ void fold_halo_sum4(const T* __restrict__ dpad, int n, int y, int x, int up, int pad, int Hu,
                                            int Wu, int Hd, int Wd, int C, int c0, float e[4]) {
  e[0] = e[1] = e[2] = e[3] = 0.f;
  for (int uy = y * up; uy < y * up + up; ++uy) {
    int er[2];
    fold_extra(uy, Hu, pad, er);
    for (int ux = x * up; ux < x * up + up; ++ux) {
      int ec[2];
      fold_extra(ux, Wu, pad, ec);
      // (extra row, own col), (own row, extra col), (extra row, extra col)
      for (int ai = -1; ai < 2; ++ai) {
        const int row = ai < 0 ? uy + pad : er[ai];
        if (row < 0) continue;
        for (int bi = -1; bi < 2; ++bi) {
          if (ai < 0 && bi < 0) continue;          // the interior copy was already added
          const int col = bi < 0 ? ux + pad : ec[bi];
          if (col < 0) continue;
          Raw4<T> h;
          float t[4];
          h.load(dpad + (((int64_t)n * Hd + row) * Wd + col) * C + c0);
          h.get(t);
          e[0] += t[0]; e[1] += t[1]; e[2] += t[2]; e[3] += t[3];
        }
      }
    }
  }
}

// backward pass 1 (see header).  One thread per source pixel x FOUR channels (8-byte loads; a warp still covers
// whole 32-byte sectors): the per-thread state is small enough for 4 resident blocks per SM without spilling --
// the 8-channel version needed 80+ registers, spilled inside the loop (3 M local-store sectors per launch) and ran
// at 2 TB/s.  Optional inputs sit behind warp-uniform branches, (n, y, x) advance without divisions, the per-
// channel scale/shift come from shared memory, and the second sum runs on the raw conv output:
// sum dz*xhat = istd * (sum dz*r - mean * sum dz) is formed once per block in double.
template <typename T, int ACT, int UP>
__global__ void __launch_bounds__(256, 4)
    bn_act_bwd_reduce_kernel(const T* __restrict__ dpad, const T* __restrict__ g1, const T* __restrict__ g2,
                             const float* __restrict__ g_nchw, const T* __restrict__ r, const float* __restrict__ sb,
                             const T* __restrict__ skip, int skip_pad, const float* __restrict__ prelu_a,
                             T* __restrict__ dz, double* __restrict__ red, double* __restrict__ dprelu, int B, int H,
                             int W, int C, int c_real, int pad) {
  constexpr int act = ACT;
  const int C4 = C >> 2;
  const int c4 = threadIdx.x % C4, lane_p = threadIdx.x / C4, ppb = 256 / C4;
  const int c0 = c4 * 4;
  const int Hu = H * UP, Wu = W * UP, Hd = Hu + 2 * pad, Wd = Wu + 2 * pad;
  const int Hs = H + 2 * skip_pad, Ws = W + 2 * skip_pad;
  const float a = prelu_a != nullptr ? *prelu_a : 0.f;
  const uint32_t pixels = (uint32_t)B * H * W;   // < 2^31, checked by the caller
  __shared__ __align__(16) float s_sc[512], s_sh[512];
  for (int c = threadIdx.x; c < C; c += 256) {
    s_sc[c] = sb != nullptr ? sb[c] : 1.f;
    s_sh[c] = sb != nullptr ? sb[C + c] : 0.f;
  }
  __syncthreads();
  float s1[4] = {0.f, 0.f, 0.f, 0.f}, s2[4] = {0.f, 0.f, 0.f, 0.f};
  float sa = 0.f;
  const int n_ok = c_real - c0;                  // channels k < n_ok of this thread are real
  uint32_t p = blockIdx.x * ppb + lane_p;
  const uint32_t step = gridDim.x * ppb;
  int x = (int)(p % (uint32_t)W), y = (int)((p / (uint32_t)W) % (uint32_t)H), n = (int)(p / ((uint32_t)W * H));
  const int dx = (int)(step % (uint32_t)W), dy = (int)((step / (uint32_t)W) % (uint32_t)H),
            dn = (int)(step / ((uint32_t)W * H));
  // Pixels within `pad` of the border also collect the mirrored copies of the reflection halo.  They are a few
  // percent of the map and take a second loop of their own, so that the halo code (and its register pressure)
  // stays out of the streaming loop.
  const bool has_halo = dpad != nullptr && pad > 0;
  auto near_border = [&](int y_, int x_) {
    const int uy0 = y_ * UP, ux0 = x_ * UP;
    return uy0 <= pad || uy0 + UP - 1 >= Hu - 1 - pad || ux0 <= pad || ux0 + UP - 1 >= Wu - 1 - pad;
  };
  auto item = [&](auto fold_tag, uint32_t p, int n, int y, int x) {
      // ---- issue every load first ----
      Raw4<T> d00, d01, d10, d11, q1, q2, rr, sk;
      const size_t off = (size_t)p * C + c0;
      rr.load(r + off);
      if (dpad != nullptr) {
        const T* dbase = dpad + (((size_t)n * Hd + y * UP + pad) * Wd + x * UP + pad) * C + c0;
        d00.load(dbase);
        if (UP == 2) {
          d01.load(dbase + C);
          d10.load(dbase + (size_t)Wd * C);
          d11.load(dbase + (size_t)Wd * C + C);
        }
      }
      if (g1 != nullptr) q1.load(g1 + off);
      if (g2 != nullptr) q2.load(g2 + off);
      if (skip != nullptr) sk.load(skip + (((size_t)n * Hs + y + skip_pad) * Ws + x + skip_pad) * C + c0);
      // ---- combine the incoming gradients ----
      float g[4] = {0.f, 0.f, 0.f, 0.f}, t[4];
      if (dpad != nullptr) {
        d00.get(g);
        if (UP == 2) {
          d01.get(t); g[0] += t[0]; g[1] += t[1]; g[2] += t[2]; g[3] += t[3];
          d10.get(t); g[0] += t[0]; g[1] += t[1]; g[2] += t[2]; g[3] += t[3];
          d11.get(t); g[0] += t[0]; g[1] += t[1]; g[2] += t[2]; g[3] += t[3];
        }
        if (decltype(fold_tag)::value) {
          float e[4];
          fold_halo_sum4<T>(dpad, n, y, x, UP, pad, Hu, Wu, Hd, Wd, C, c0, e);
          g[0] += e[0]; g[1] += e[1]; g[2] += e[2]; g[3] += e[3];
        }
      }
      if (g1 != nullptr) { q1.get(t); g[0] += t[0]; g[1] += t[1]; g[2] += t[2]; g[3] += t[3]; }
      if (g2 != nullptr) { q2.get(t); g[0] += t[0]; g[1] += t[1]; g[2] += t[2]; g[3] += t[3]; }
      if (g_nchw != nullptr) {
  #pragma unroll
        for (int k = 0; k < 4; ++k)
          if (k < n_ok) g[k] += g_nchw[(((size_t)n * c_real + c0 + k) * H + y) * W + x];
      }
      // ---- through the activation, sums ----
      float rv[4], o[4];
      rr.get(rv);
      if (act != P2P_ACT_NONE) {
        float skv[4] = {0.f, 0.f, 0.f, 0.f};
        if (skip != nullptr) sk.get(skv);
        const float4 sc = *reinterpret_cast<const float4*>(s_sc + c0), sf = *reinterpret_cast<const float4*>(s_sh + c0);
        const float scv[4] = {sc.x, sc.y, sc.z, sc.w}, sfv[4] = {sf.x, sf.y, sf.z, sf.w};
  #pragma unroll
        for (int k = 0; k < 4; ++k) {
          const float z = fmaf(rv[k], scv[k], sfv[k]) + skv[k];
          if (act == P2P_ACT_PRELU && z < 0.f && k < n_ok) sa += g[k] * z;
          g[k] *= act_bwd(z, act, a);
        }
      }
  #pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float d = k < n_ok ? g[k] : 0.f;
        o[k] = d;
        const float dq = round_as(d, dz);
        s1[k] += dq;
        s2[k] = fmaf(dq, rv[k], s2[k]);
      }
      Raw4<T>::store(dz + off, o);
  };
  auto advance = [&](int& n_, int& y_, int& x_) {
    x_ += dx;
    if (x_ >= W) { x_ -= W; ++y_; }
    y_ += dy;
    if (y_ >= H) { y_ -= H; ++n_; }
    n_ += dn;
  };
  const uint32_t p0 = p;
  const int x0 = x, y0 = y, n0 = n;
  for (; p < pixels; p += step) {
    if (!(has_halo && near_border(y, x))) item(std::false_type{}, p, n, y, x);
    advance(n, y, x);
  }
  if (has_halo) {
    p = p0; x = x0; y = y0; n = n0;
    for (; p < pixels; p += step) {
      if (near_border(y, x)) item(std::true_type{}, p, n, y, x);
      advance(n, y, x);
    }
  }
  if (red != nullptr) {
    __shared__ float sh[2][256 * 5];           // [which][thread * 5 + k]: conflict-free writes and column reads
#pragma unroll
    for (int k = 0; k < 4; ++k) { sh[0][threadIdx.x * 5 + k] = s1[k]; sh[1][threadIdx.x * 5 + k] = s2[k]; }
    __syncthreads();
    for (int c = threadIdx.x; c < C; c += 256) {
      double a1 = 0.0, a2 = 0.0;
      for (int l = 0; l < ppb; ++l) {
        const int idx = (l * C4 + (c >> 2)) * 5 + (c & 3);
        a1 += (double)sh[0][idx];
        a2 += (double)sh[1][idx];
      }
      if (c < c_real) {
        const double mean = sb != nullptr ? (double)sb[2 * C + c] : 0.0;
        const double istd = sb != nullptr ? (double)sb[3 * C + c] : 0.0;
        atomicAdd(red + c, a1);
        atomicAdd(red + C + c, istd * (a2 - mean * a1));
      }
    }
  }
  if (dprelu != nullptr && act == P2P_ACT_PRELU) {
    const float tot = block_sum<float>(sa);
    if (threadIdx.x == 0) atomicAdd(dprelu, (double)tot);
  }
}

// backward pass 2:  dr = scale * (dz - mean(dz) - xhat * mean(dz * xhat))  =  A*dz + Bx*(r - mean) + K  per channel.
// Four channels per thread, the per-channel coefficients in shared memory (same treatment as pass 1: 32 registers
// of per-channel state per thread became none).
template <typename T>
__global__ void __launch_bounds__(256, 4)
    bn_bwd_apply_kernel(const T* __restrict__ dz, const T* __restrict__ r, const float* __restrict__ sb,
                        const double* __restrict__ red, T* __restrict__ dr, float* __restrict__ dbias,
                        float* __restrict__ dgamma, float* __restrict__ dbeta, int64_t pixels, int C, int c_real) {
  __shared__ __align__(16) float s_a[512], s_bx[512], s_mean[512], s_k[512];
  __shared__ float sh[256 * 5];
  const double inv = 1.0 / (double)pixels;
  for (int c = threadIdx.x; c < C; c += 256) {
    const bool real = c < c_real;
    const float scale = real ? sb[c] : 0.f, istd = sb[3 * C + c];
    const float m1 = real ? (float)(red[c] * inv) : 0.f, m2 = real ? (float)(red[C + c] * inv) : 0.f;
    s_a[c] = scale;
    s_bx[c] = -scale * istd * m2;
    s_mean[c] = sb[2 * C + c];
    s_k[c] = -scale * m1;
    if (blockIdx.x == 0 && real && dgamma != nullptr && dbeta != nullptr) {   // d gamma = sum dz * xhat, d beta = sum dz
      dbeta[c] += (float)red[c];
      dgamma[c] += (float)red[C + c];
    }
  }
  __syncthreads();
  const int C4 = C >> 2;
  const int c4 = threadIdx.x % C4, lane_p = threadIdx.x / C4, ppb = 256 / C4;
  const int c0 = c4 * 4;
  float sd[4] = {0.f, 0.f, 0.f, 0.f};
  for (int64_t p = (int64_t)blockIdx.x * ppb + lane_p; p < pixels; p += (int64_t)gridDim.x * ppb) {
    Raw4<T> dd, rr;
    dd.load(dz + (size_t)p * C + c0);
    rr.load(r + (size_t)p * C + c0);
    const float4 a = *reinterpret_cast<const float4*>(s_a + c0), bx = *reinterpret_cast<const float4*>(s_bx + c0);
    const float4 mu = *reinterpret_cast<const float4*>(s_mean + c0), kk = *reinterpret_cast<const float4*>(s_k + c0);
    float d[4], rv[4];
    dd.get(d);
    rr.get(rv);
    d[0] = fmaf(a.x, d[0], fmaf(bx.x, rv[0] - mu.x, kk.x));
    d[1] = fmaf(a.y, d[1], fmaf(bx.y, rv[1] - mu.y, kk.y));
    d[2] = fmaf(a.z, d[2], fmaf(bx.z, rv[2] - mu.z, kk.z));
    d[3] = fmaf(a.w, d[3], fmaf(bx.w, rv[3] - mu.w, kk.w));
    sd[0] += d[0]; sd[1] += d[1]; sd[2] += d[2]; sd[3] += d[3];
    Raw4<T>::store(dr + (size_t)p * C + c0, d);
  }
  if (dbias != nullptr) {
#pragma unroll
    for (int k = 0; k < 4; ++k) sh[threadIdx.x * 5 + k] = sd[k];
    __syncthreads();
    for (int c = threadIdx.x; c < c_real; c += 256) {
      float acc = 0.f;
      for (int l = 0; l < ppb; ++l) acc += sh[(l * C4 + (c >> 2)) * 5 + (c & 3)];
      atomicAdd(dbias + c, acc);
    }
  }
}

__global__ void prelu_grad_finish_kernel(const double* d, float* g) { *g += (float)(*d); }

// =================================================================================================
// discriminator input pyramid
// =================================================================================================
__global__ void d_input_prep_kernel(const float* __restrict__ a, int ca, const float* __restrict__ b, int cb,
                                    float* __restrict__ dst, int B, int H, int W) {
  const int64_t total = (int64_t)B * H * W;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t hw = (int64_t)H * W;
    const int n = (int)(i / hw);
    const int64_t rem = i - (int64_t)n * hw;
    float v[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      float t = 0.f;
      if (k < ca) t = a[((int64_t)n * ca + k) * hw + rem];
      else if (b != nullptr && k < ca + cb) t = b[((int64_t)n * cb + (k - ca)) * hw + rem];
      v[k] = t;
    }
    st8(dst + i * 8, v);
  }
}

// AvgPool2d(3, 2, 1, count_include_pad=False), NHWC fp32 with C == 8
__global__ void avgpool_fwd_kernel(const float* __restrict__ in, float* __restrict__ out, int B, int H, int W, int Ho,
                                   int Wo) {
  const int64_t total = (int64_t)B * Ho * Wo;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int ox = (int)(i % Wo);
    const int oy = (int)((i / Wo) % Ho);
    const int n = (int)(i / ((int64_t)Wo * Ho));
    float acc[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[k] = 0.f;
    int cnt = 0;
    for (int ky = 0; ky < 3; ++ky) {
      const int y = oy * 2 - 1 + ky;
      if (y < 0 || y >= H) continue;
      for (int kx = 0; kx < 3; ++kx) {
        const int x = ox * 2 - 1 + kx;
        if (x < 0 || x >= W) continue;
        float v[8];
        ld8(in + (((int64_t)n * H + y) * W + x) * 8, v);
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[k] += v[k];
        ++cnt;
      }
    }
    const float invc = 1.f / (float)cnt;
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[k] *= invc;
    st8(out + i * 8, acc);
  }
}

__device__ __forceinline__ int pool_count(int o, int n) {  // valid taps of output o along one axis
  int c = 0;
  for (int k = 0; k < 3; ++k) {
    const int y = o * 2 - 1 + k;
    c += (y >= 0 && y < n);
  }
  return c;
}

__global__ void avgpool_bwd_add_kernel(const float* __restrict__ dout, float* __restrict__ din, int B, int H, int W,
                                       int Ho, int Wo) {
  const int64_t total = (int64_t)B * H * W;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int x = (int)(i % W);
    const int y = (int)((i / W) % H);
    const int n = (int)(i / ((int64_t)W * H));
    float acc[8];
    ld8(din + i * 8, acc);
    // outputs oy with oy*2-1 <= y <= oy*2+1
    for (int oy = (y) / 2; oy <= (y + 1) / 2; ++oy) {
      if (oy < 0 || oy >= Ho || y < oy * 2 - 1 || y > oy * 2 + 1) continue;
      const int cy = pool_count(oy, H);
      for (int ox = (x) / 2; ox <= (x + 1) / 2; ++ox) {
        if (ox < 0 || ox >= Wo || x < ox * 2 - 1 || x > ox * 2 + 1) continue;
        const float wgt = 1.f / (float)(cy * pool_count(ox, W));
        float v[8];
        ld8(dout + (((int64_t)n * Ho + oy) * Wo + ox) * 8, v);
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[k] += v[k] * wgt;
      }
    }
    st8(din + i * 8, acc);
  }
}

template <typename T>
__global__ void cast_pad_channels_kernel(const float* __restrict__ in, T* __restrict__ out, int64_t pixels, int Cs,
                                         int Cd) {
  const int D8 = Cd >> 3;
  const int64_t total = pixels * D8;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c8 = (int)(i % D8);
    const int64_t p = i / D8;
    float v[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = 0.f;
    if (c8 * 8 < Cs) ld8(in + p * Cs + c8 * 8, v);
    st8(out + i * 8, v);
  }
}

template <typename T>
__global__ void uncast_add_kernel(const T* __restrict__ in, float* __restrict__ out, int64_t pixels, int Cs, int Cd) {
  const int S8 = Cs >> 3;
  const int64_t total = pixels * S8;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c8 = (int)(i % S8);
    const int64_t p = i / S8;
    float v[8], o[8];
    ld8(in + p * Cd + c8 * 8, v);
    ld8(out + i * 8, o);
#pragma unroll
    for (int k = 0; k < 8; ++k) o[k] += v[k];
    st8(out + i * 8, o);
  }
}

__global__ void d_input_grad_kernel(const float* __restrict__ dx, float* __restrict__ db, int ca, int cb, int B, int H,
                                    int W) {
  const int64_t hw = (int64_t)H * W, total = (int64_t)B * cb * hw;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t rem = i % hw;
    const int c = (int)((i / hw) % cb);
    const int n = (int)(i / (hw * cb));
    db[i] = dx[((int64_t)n * hw + rem) * 8 + ca + c];
  }
}

// dz = (ga + gb) * act'(out), plus per-channel sum for the conv bias gradient
template <typename T>
__global__ void __launch_bounds__(256)
    act_bwd_kernel(const T* __restrict__ out, const T* __restrict__ ga, const T* __restrict__ gb, T* __restrict__ dz,
                   float* __restrict__ dbias, int64_t pixels, int C, int c_real, int act) {
  const int C8 = C >> 3;
  const int c8 = threadIdx.x % C8, lane_p = threadIdx.x / C8, ppb = 256 / C8;
  float sd[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) sd[k] = 0.f;
  for (int64_t p = (int64_t)blockIdx.x * ppb + lane_p; p < pixels; p += (int64_t)gridDim.x * ppb) {
    float g[8], t[8], o[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) g[k] = 0.f;
    if (ga != nullptr) {
      ld8(ga + (size_t)p * C + c8 * 8, t);
#pragma unroll
      for (int k = 0; k < 8; ++k) g[k] += t[k];
    }
    if (gb != nullptr) {
      ld8(gb + (size_t)p * C + c8 * 8, t);
#pragma unroll
      for (int k = 0; k < 8; ++k) g[k] += t[k];
    }
    if (act == P2P_ACT_LRELU || act == P2P_ACT_RELU) {
      const float neg = act == P2P_ACT_LRELU ? 0.2f : 0.f;
      ld8(out + (size_t)p * C + c8 * 8, o);
#pragma unroll
      for (int k = 0; k < 8; ++k) g[k] *= (o[k] > 0.f ? 1.f : neg);
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      if (c8 * 8 + k >= c_real) g[k] = 0.f;
      sd[k] += round_as(g[k], dz);
    }
    st8(dz + (size_t)p * C + c8 * 8, g);
  }
  if (dbias != nullptr) {
    __shared__ float sh[8][256];
#pragma unroll
    for (int k = 0; k < 8; ++k) sh[k][threadIdx.x] = sd[k];
    __syncthreads();
    for (int c = threadIdx.x; c < c_real; c += 256) {
      float acc = 0.f;
      for (int l = 0; l < ppb; ++l) acc += sh[c & 7][l * C8 + (c >> 3)];
      atomicAdd(dbias + c, acc);
    }
  }
}

template <typename T>
__global__ void nhwc_to_nchw_kernel(const T* __restrict__ in, float* __restrict__ out, int B, int H, int W, int C,
                                    int c_real) {
  const int64_t hw = (int64_t)H * W, total = (int64_t)B * c_real * hw;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t rem = i % hw;
    const int c = (int)((i / hw) % c_real);
    const int n = (int)(i / (hw * c_real));
    out[i] = ldf(in + ((int64_t)n * hw + rem) * C + c);
  }
}

template <typename T>
__global__ void nchw_to_nhwc_kernel(const float* __restrict__ in, T* __restrict__ out, int B, int H, int W, int C,
                                    int c_real) {
  const int C8 = C >> 3;
  const int64_t hw = (int64_t)H * W, total = (int64_t)B * hw * C8;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c8 = (int)(i % C8);
    const int64_t p = i / C8;
    const int n = (int)(p / hw);
    const int64_t rem = p - (int64_t)n * hw;
    float v[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const int c = c8 * 8 + k;
      v[k] = c < c_real ? in[((int64_t)n * c_real + c) * hw + rem] : 0.f;
    }
    st8(out + i * 8, v);
  }
}

static inline int ew_grid(int64_t total, int block = 256) {
  int64_t g = (total + block - 1) / block;
  const int64_t cap = (int64_t)kSMs * 16;
  return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}
// kernels that end in per-channel atomics: 4 resident blocks per SM, so each address sees <= 592 adds
static inline int red_ew_grid(int64_t pixels, int ppb) {
  int64_t g = (pixels + ppb - 1) / ppb;
  const int64_t cap = (int64_t)kSMs * 4;
  return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

// =================================================================================================
// Convolution with one output channel as "1x1 convolution with one output column per filter tap" + this gather:
//   out[n, oy, ox, 0] = bias + sum_{ky,kx} t[n, oy + ky - pad, ox + kx - pad, ky*kw + kx]      (zero outside)
// t = the input map times the [taps x Ci] filter, one value per (input pixel, tap).  Channels 1..15 of out are 0.
// =================================================================================================
template <typename T>
__global__ void __launch_bounds__(256) tap_gather_sum_kernel(const T* __restrict__ t, const float* __restrict__ bias,
                                                             T* __restrict__ out, int N, int Hi, int Wi, int Ho, int Wo,
                                                             int kh, int kw, int pad, int act) {
  const int64_t total = (int64_t)N * Ho * Wo;
  const float b0 = bias != nullptr ? bias[0] : 0.f;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int ox = (int)(i % Wo), oy = (int)((i / Wo) % Ho), n = (int)(i / ((int64_t)Wo * Ho));
    float acc = b0;
    for (int ky = 0; ky < kh; ++ky) {
      const int iy = oy + ky - pad;
      if (iy < 0 || iy >= Hi) continue;
      for (int kx = 0; kx < kw; ++kx) {
        const int ix = ox + kx - pad;
        if (ix < 0 || ix >= Wi) continue;
        acc += ldf(t + (((size_t)n * Hi + iy) * Wi + ix) * 16 + ky * kw + kx);
      }
    }
    if (act == P2P_ACT_LRELU) acc = fmaxf(acc, 0.2f * acc);
    float v[8] = {acc, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f}, z[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    st8(out + (size_t)i * 16, v);
    st8(out + (size_t)i * 16 + 8, z);
  }
}

// the data-gradient counterpart: a[n, iy, ix, ky*kw + kx] = dy[n, iy + pad - ky, ix + pad - kx, 0]  (zero outside);
// a 1x1 convolution of `a` with the [Ci x taps] filter matrix is then the gradient w.r.t. the layer input
template <typename T>
__global__ void __launch_bounds__(256) tap_scatter_kernel(const T* __restrict__ dy, T* __restrict__ a, int N, int Hi,
                                                          int Wi, int Ho, int Wo, int kh, int kw, int pad) {
  const int64_t total = (int64_t)N * Hi * Wi;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int ix = (int)(i % Wi), iy = (int)((i / Wi) % Hi), n = (int)(i / ((int64_t)Wi * Hi));
    float v[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) v[k] = 0.f;
#pragma unroll
    for (int k = 0; k < 16; ++k) {
      const int ky = k / kw, kx = k - ky * kw;
      const int oy = iy + pad - ky, ox = ix + pad - kx;
      if (k < kh * kw && oy >= 0 && oy < Ho && ox >= 0 && ox < Wo) v[k] = ldf(dy + (((size_t)n * Ho + oy) * Wo + ox) * 16);
    }
    st8(a + (size_t)i * 16, v);
    st8(a + (size_t)i * 16 + 8, v + 8);
  }
}

// =================================================================================================
// Wide filters with very few output channels (the generator's last layer, 9x9, 32 -> 3): the convolution runs as a
// kh x 1 convolution with kw * co "fat" output channels (one MMA column per (kx, co) pair instead of 16 columns
// per tap), and these two kernels are the horizontal part:
//   gather:   out[n, y, x, co] = bias[co] + sum_kx t[n, y, x + kx, kx * co_real + co]
//   scatter:  a[n, y, xa, kx * co_real + co] = dy[n, y, xa - kx, co]                    (the adjoint, for dgrad)
// =================================================================================================
template <typename T>
__global__ void __launch_bounds__(256) kx_gather_sum_kernel(const T* __restrict__ t, const float* __restrict__ bias,
                                                            T* __restrict__ out, int H, int Wt, int Wo, int kw,
                                                            int co_real, int Ct, int Co) {
  extern __shared__ __align__(16) uint32_t s_t[];          // [(256 + kw - 1) pixels][Ct/2 + 1 words]
  const int pitch = Ct / 2 + 1;
  const int x0 = blockIdx.x * 256, y = blockIdx.y, n = blockIdx.z;
  const int npix = min(256, Wo - x0) + kw - 1;
  const uint32_t* src = reinterpret_cast<const uint32_t*>(t + (((size_t)n * H + y) * Wt + x0) * Ct);   // bf16 pairs
  if (sizeof(T) == 2) {
    for (int e = threadIdx.x; e < npix * (Ct / 2); e += 256) s_t[(e / (Ct / 2)) * pitch + e % (Ct / 2)] = src[e];
  }
  __syncthreads();
  const int x = x0 + threadIdx.x;
  if (x >= Wo) return;
  float acc[8];
#pragma unroll
  for (int c = 0; c < 8; ++c) acc[c] = (c < co_real && bias != nullptr) ? bias[c] : 0.f;
  for (int kx = 0; kx < kw; ++kx) {
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      if (c >= co_real) break;
      const int ch = kx * co_real + c;
      float v;
      if (sizeof(T) == 2) {
        const uint32_t w = s_t[(threadIdx.x + kx) * pitch + (ch >> 1)];
        v = __uint_as_float((ch & 1) ? (w & 0xffff0000u) : (w << 16));
      } else {
        v = ldf(t + (((size_t)n * H + y) * Wt + x + kx) * Ct + ch);
      }
      acc[c] += v;
    }
  }
  T* o = out + (((size_t)n * H + y) * Wo + x) * Co;
  for (int c0 = 0; c0 < Co; c0 += 8) {
    float v[8];
#pragma unroll
    for (int c = 0; c < 8; ++c) v[c] = (c0 == 0 && c < co_real) ? acc[c] : 0.f;
    st8(o + c0, v);
  }
}

template <typename T>
__global__ void __launch_bounds__(256) kx_scatter_kernel(const T* __restrict__ dy, T* __restrict__ a, int N, int H, int Wo,
                                                         int Wa, int kw, int co_real, int Cs, int Ca) {
  const int64_t total = (int64_t)N * H * Wa * (Ca / 8);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c8 = (int)(i % (Ca / 8));
    int64_t q = i / (Ca / 8);
    const int xa = (int)(q % Wa); q /= Wa;               // q = n * H + y
    float v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int ch = c8 * 8 + j, kx = ch / co_real, co = ch - kx * co_real;
      const int x = xa - kx;
      v[j] = (kx < kw && x >= 0 && x < Wo) ? ldf(dy + ((size_t)q * Wo + x) * Cs + co) : 0.f;
    }
    st8(a + (size_t)i * 8, v);
  }
}

// =================================================================================================
// Tail of the compression front end (CompressionNetwork.forward, networks.py:226-236, + compress, train.py:297):
// PixelShuffle(2) of the 12-channel map, L2-normalise over the 3 output channels, add the input image, and
// optionally quantise to 2^bits - 1 levels (generate_dataset.py:29-34).  One thread per low-resolution pixel.
// =================================================================================================
template <typename T>
__global__ void __launch_bounds__(256) c_output_kernel(const T* __restrict__ res, const float* __restrict__ x,
                                                       float* __restrict__ y, float* __restrict__ comp, int B, int H,
                                                       int W, float levels) {
  const int h2 = H >> 1, w2 = W >> 1;
  const int64_t total = (int64_t)B * h2 * w2;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int j = (int)(i % w2), ii = (int)((i / w2) % h2), n = (int)(i / ((int64_t)w2 * h2));
    float v[16];
    ld8(res + (size_t)i * 16, v);
    ld8(res + (size_t)i * 16 + 8, v + 8);
#pragma unroll
    for (int s = 0; s < 4; ++s) {                     // sub-pixel (dy, dx) = (s >> 1, s & 1); channel c reads 4c + s
      const float a = v[s], b = v[4 + s], c = v[8 + s];
      const float inv = 1.f / fmaxf(sqrtf(a * a + b * b + c * c), 1e-12f);      // F.normalize(p=2, dim=1, eps=1e-12)
      const int yy = 2 * ii + (s >> 1), xx = 2 * j + (s & 1);
      const float r3[3] = {a * inv, b * inv, c * inv};
#pragma unroll
      for (int ch = 0; ch < 3; ++ch) {
        const size_t o = (((size_t)n * 3 + ch) * H + yy) * W + xx;
        const float out = x[o] + r3[ch];
        if (y != nullptr) y[o] = out;
        if (comp != nullptr) comp[o] = rintf(fminf(fmaxf(out, 0.f), 1.f) * levels) / levels;
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------------
// Zero-fill with 16-byte stores.  The weight-gradient accumulators (split-K atomics) must start at zero: 33 MB per
// discriminator backward pass, 11.5 MB for the generator; torch.zeros() on a tensor of that size launches a
// byte-granular fill kernel that ncu times at ~70 us (0.47 TB/s).
// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) zero_fill_kernel(uint4* __restrict__ p, int64_t n16, uint8_t* __restrict__ tail,
                                                        int n_tail) {
  const uint4 z = make_uint4(0u, 0u, 0u, 0u);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (int64_t)gridDim.x * blockDim.x) p[i] = z;
  if (blockIdx.x == 0 && (int)threadIdx.x < n_tail) tail[threadIdx.x] = 0;
}

// ---------------------------------------------------------------------------------------------------
// 8-bit image batch -> the fp32 tensors the iteration consumes (dataset.py:19-20,30-40: ToTensor = x / 255,
// Normalize(0.5, 0.5) = (t - 0.5) / 0.5), same fp32 operations in the same order as torchvision, and optionally
// the 3-bit quantised generator input compress(b * 0.5 + 0.5, bits) (generate_dataset.py:29-34) in the same pass.
// 16 pixels-channels per thread: one 16-byte load, four 16-byte stores per output.
// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ float decode_u8(uint32_t v) { return __fdiv_rn(__fdiv_rn((float)v, 255.f) - 0.5f, 0.5f); }
__device__ __forceinline__ float quantise_unit(float real, float q) {
  const float t = fminf(fmaxf(real * 0.5f + 0.5f, 0.f), 1.f) * q;
  return __fdiv_rn(rintf(t), q);          // torch.round is round-half-to-even, like rintf
}
__global__ void __launch_bounds__(256) decode_u8_kernel(const uint8_t* __restrict__ src, float* __restrict__ dst,
                                                        float* __restrict__ quant, int64_t n, float q) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x * 16;
  for (int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 16; i < n; i += stride) {
    if (i + 16 <= n) {
      const uint4 u = *reinterpret_cast<const uint4*>(src + i);
      const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        float4 f;
        f.x = decode_u8(w[k] & 0xffu); f.y = decode_u8((w[k] >> 8) & 0xffu);
        f.z = decode_u8((w[k] >> 16) & 0xffu); f.w = decode_u8(w[k] >> 24);
        *reinterpret_cast<float4*>(dst + i + 4 * k) = f;
        if (quant != nullptr)
          *reinterpret_cast<float4*>(quant + i + 4 * k) =
              make_float4(quantise_unit(f.x, q), quantise_unit(f.y, q), quantise_unit(f.z, q), quantise_unit(f.w, q));
      }
    } else {
      for (int64_t j = i; j < n; ++j) {
        const float f = decode_u8(src[j]);
        dst[j] = f;
        if (quant != nullptr) quant[j] = quantise_unit(f, q);
      }
    }
  }
}


// ---------------------------------------------------------------------------------------------------
// MaxPool2d(kernel 2, stride 2) on stored NHWC maps (torchvision's vgg19.features[4, 9, 18, 27], reference
// networks.py:32-50) and its gradient: the whole gradient goes to the first maximum of the window in scan order
// (PyTorch keeps the index of the first element that compares greater).  8 channels per thread.
// ---------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256)
    maxpool2_fwd_kernel(const T* __restrict__ in, T* __restrict__ out, int N, int H, int W, int C) {
  const int Ho = H >> 1, Wo = W >> 1, C8 = C >> 3;
  const int64_t total = (int64_t)N * Ho * Wo * C8;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c8 = (int)(i % C8);
    int64_t r = i / C8;
    const int x = (int)(r % Wo);
    r /= Wo;
    const int y = (int)(r % Ho), n = (int)(r / Ho);
    const T* src = in + (((int64_t)n * H + 2 * y) * W + 2 * x) * C + c8 * 8;
    float m[8], t[8];
    ld8(src, m);
    ld8(src + C, t);
#pragma unroll
    for (int k = 0; k < 8; ++k) m[k] = t[k] > m[k] ? t[k] : m[k];
    ld8(src + (int64_t)W * C, t);
#pragma unroll
    for (int k = 0; k < 8; ++k) m[k] = t[k] > m[k] ? t[k] : m[k];
    ld8(src + (int64_t)W * C + C, t);
#pragma unroll
    for (int k = 0; k < 8; ++k) m[k] = t[k] > m[k] ? t[k] : m[k];
    st8(out + i * 8, m);
  }
}

template <typename T>
__global__ void __launch_bounds__(256)
    maxpool2_bwd_kernel(const T* __restrict__ in, const T* __restrict__ dout, T* __restrict__ din, int N, int H, int W,
                        int C) {
  const int Ho = H >> 1, Wo = W >> 1, C8 = C >> 3;
  const int64_t total = (int64_t)N * Ho * Wo * C8;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c8 = (int)(i % C8);
    int64_t r = i / C8;
    const int x = (int)(r % Wo);
    r /= Wo;
    const int y = (int)(r % Ho), n = (int)(r / Ho);
    const int64_t base = (((int64_t)n * H + 2 * y) * W + 2 * x) * C + c8 * 8;
    const int64_t offs[4] = {0, C, (int64_t)W * C, (int64_t)W * C + C};
    float v[4][8], g[8], z[8];
    int arg[8];
#pragma unroll
    for (int q = 0; q < 4; ++q) ld8(in + base + offs[q], v[q]);
    ld8(dout + i * 8, g);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      arg[k] = 0;
      float m = v[0][k];
#pragma unroll
      for (int q = 1; q < 4; ++q)
        if (v[q][k] > m) { m = v[q][k]; arg[k] = q; }
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
#pragma unroll
      for (int k = 0; k < 8; ++k) z[k] = arg[k] == q ? g[k] : 0.f;
      st8(din + base + offs[q], z);
    }
  }
}

}  // namespace p2p

using namespace p2p;

extern "C" {


int p2p_g_input_prep(const float* x, void* dst, int B, int H, int W, int dtype, void* stream) {
  P2P_REQUIRE(x && dst && B > 0 && H > 0 && W > 0, "null pointer or empty shape");
  P2P_REQUIRE(H % 4 == 0 && W % 4 == 0 && H >= 8 && W >= 8, "H and W must be multiples of 4 and >= 8");
  const int64_t total = (int64_t)B * (H + 8) * (W + 8) * 2;
  P2P_DISPATCH(dtype, g_input_prep_kernel<T><<<ew_grid(total), 256, 0, as_stream(stream)>>>(x, (T*)dst, B, H, W););
  P2P_LAUNCH_CHECK();
}

int p2p_channel_stats(const void* r, double* stats, int64_t pixels, int C, int dtype, void* stream) {
  P2P_REQUIRE(r && stats && pixels > 0, "null pointer or empty map");
  P2P_REQUIRE(C % 16 == 0 && C <= 512, "C must be a multiple of 16, <= 512");
  const int ppb = 256 / (C / 8);
  P2P_DISPATCH(dtype, channel_stats_kernel<T><<<red_ew_grid(pixels, ppb), 256, 0, as_stream(stream)>>>((const T*)r, stats,
                                                                                                pixels, C););
  P2P_LAUNCH_CHECK();
}

int p2p_bn_finalize(const double* stats, int64_t count, const float* gamma, const float* beta, float* running_mean,
                    float* running_var, int64_t* num_batches_tracked, float* sb, int c_real, int C, int training,
                    void* stream) {
  P2P_REQUIRE(gamma && beta && sb, "null pointer");
  P2P_REQUIRE(training ? (stats != nullptr && count > 0) : (running_mean && running_var), "missing statistics");
  bn_finalize_kernel<<<cdiv(C, 128), 128, 0, as_stream(stream)>>>(stats, count, gamma, beta, running_mean, running_var,
                                                                 num_batches_tracked, sb, c_real, C, training);
  P2P_LAUNCH_CHECK();
}

int p2p_bn_act_store(const void* r, const float* sb, const void* skip, int skip_pad, const float* prelu_a, void* dst,
                     float* out_nchw, int B, int H, int W, int C, int c_real, int act, int up, int pad, int dtype,
                     void* stream) {
  P2P_REQUIRE(r && (dst || out_nchw), "null pointer");
  P2P_REQUIRE(C % 16 == 0 && (up == 1 || up == 2) && pad >= 0, "bad geometry");
  P2P_REQUIRE(pad < H * up && pad < W * up, "reflection pad must be smaller than the image");
  P2P_REQUIRE(act != P2P_ACT_PRELU || prelu_a, "PReLU needs its slope");
  if (out_nchw != nullptr) {
    P2P_REQUIRE(skip == nullptr && up == 1, "NCHW output takes no skip / upsample");
    const int64_t total = (int64_t)B * c_real * H * W;
    P2P_DISPATCH(dtype, bn_act_store_nchw_kernel<T><<<ew_grid(total), 256, 0, as_stream(stream)>>>(
                            (const T*)r, sb, prelu_a, out_nchw, B, H, W, C, c_real, act););
    P2P_LAUNCH_CHECK();
  }
  const int64_t total = (int64_t)B * (H * up + 2 * pad) * (W * up + 2 * pad) * (C / 8);
  P2P_REQUIRE(total < (1LL << 31), "tensor too large");
#define P2P_STORE(ACT_)                                                                            \
  P2P_DISPATCH(dtype, bn_act_store_kernel<T, ACT_><<<ew_grid(total), 256, 0, as_stream(stream)>>>( \
                          (const T*)r, sb, (const T*)skip, skip_pad, prelu_a, (T*)dst, B, H, W, C, up, pad););
  switch (act) {
    case P2P_ACT_NONE: P2P_STORE(P2P_ACT_NONE) break;
    case P2P_ACT_RELU: P2P_STORE(P2P_ACT_RELU) break;
    case P2P_ACT_PRELU: P2P_STORE(P2P_ACT_PRELU) break;
    case P2P_ACT_LRELU: P2P_STORE(P2P_ACT_LRELU) break;
    case P2P_ACT_TANH: P2P_STORE(P2P_ACT_TANH) break;
    default: P2P_REQUIRE(false, "unknown activation");
  }
#undef P2P_STORE
  P2P_LAUNCH_CHECK();
}

int p2p_bn_finalize_act_store(const double* stats, int64_t count, const float* gamma, const float* beta,
                              float* running_mean, float* running_var, int64_t* num_batches_tracked, float* sb,
                              int training, const void* r, const void* skip, int skip_pad, const float* prelu_a, void* dst,
                              int B, int H, int W, int C, int c_real, int act, int up, int pad, int dtype, void* stream) {
  P2P_REQUIRE(gamma && beta && sb && r && dst, "null pointer");
  P2P_REQUIRE(training ? (stats != nullptr && count > 0) : (running_mean && running_var), "missing statistics");
  P2P_REQUIRE(C % 16 == 0 && C <= 512 && (up == 1 || up == 2) && pad >= 0, "bad geometry");
  P2P_REQUIRE(pad < H * up && pad < W * up, "reflection pad must be smaller than the image");
  P2P_REQUIRE(act != P2P_ACT_PRELU || prelu_a, "PReLU needs its slope");
  const int64_t total = (int64_t)B * (H * up + 2 * pad) * (W * up + 2 * pad) * (C / 8);
  P2P_REQUIRE(total < (1LL << 31), "tensor too large");
#define P2P_FSTORE(ACT_)                                                                                     \
  P2P_DISPATCH(dtype, bn_finalize_act_store_kernel<T, ACT_><<<ew_grid(total), 256, 0, as_stream(stream)>>>( \
                          stats, count, gamma, beta, running_mean, running_var, num_batches_tracked, sb, c_real,   \
                          training, (const T*)r, (const T*)skip, skip_pad, prelu_a, (T*)dst, B, H, W, C, up, pad););
  switch (act) {
    case P2P_ACT_NONE: P2P_FSTORE(P2P_ACT_NONE) break;
    case P2P_ACT_RELU: P2P_FSTORE(P2P_ACT_RELU) break;
    case P2P_ACT_PRELU: P2P_FSTORE(P2P_ACT_PRELU) break;
    case P2P_ACT_LRELU: P2P_FSTORE(P2P_ACT_LRELU) break;
    case P2P_ACT_TANH: P2P_FSTORE(P2P_ACT_TANH) break;
    default: P2P_REQUIRE(false, "unknown activation");
  }
#undef P2P_FSTORE
  P2P_LAUNCH_CHECK();
}

int p2p_bn_act_bwd_reduce(const void* dpad, const void* g1, const void* g2, const float* g_nchw, const void* r,
                          const float* sb, const void* skip, int skip_pad, const float* prelu_a, void* dz, double* red,
                          double* dprelu, int B, int H, int W, int C, int c_real, int act, int up, int pad, int dtype,
                          void* stream) {
  P2P_REQUIRE(r && dz, "null pointer");
  P2P_REQUIRE(dpad || g1 || g2 || g_nchw, "no incoming gradient");
  P2P_REQUIRE(C % 16 == 0 && C <= 512 && (up == 1 || up == 2), "bad geometry");
  P2P_REQUIRE(act != P2P_ACT_PRELU || prelu_a, "PReLU needs its slope");
  P2P_REQUIRE(256 % (C / 4) == 0, "C must be 16, 32, 64, 128, 256 or 512");
  const int ppb = 256 / (C / 4);
  const int64_t pixels = (int64_t)B * H * W;
  const int grid = red_ew_grid(pixels, ppb);       // 4 resident blocks per SM
#define P2P_BWD_REDUCE_U(ACT_, UP_)                                                                              \
  P2P_DISPATCH(dtype, bn_act_bwd_reduce_kernel<T, ACT_, UP_><<<grid, 256, 0, as_stream(stream)>>>(               \
                          (const T*)dpad, (const T*)g1, (const T*)g2, g_nchw, (const T*)r, sb, (const T*)skip, skip_pad, \
                          prelu_a, (T*)dz, red, dprelu, B, H, W, C, c_real, pad););
#define P2P_BWD_REDUCE(ACT_)                                         \
  if (up == 2) { P2P_BWD_REDUCE_U(ACT_, 2) }                         \
  else { P2P_BWD_REDUCE_U(ACT_, 1) }
  switch (act) {
    case P2P_ACT_NONE: P2P_BWD_REDUCE(P2P_ACT_NONE) break;
    case P2P_ACT_RELU: P2P_BWD_REDUCE(P2P_ACT_RELU) break;
    case P2P_ACT_PRELU: P2P_BWD_REDUCE(P2P_ACT_PRELU) break;
    case P2P_ACT_LRELU: P2P_BWD_REDUCE(P2P_ACT_LRELU) break;
    case P2P_ACT_TANH: P2P_BWD_REDUCE(P2P_ACT_TANH) break;
    default: P2P_REQUIRE(false, "unknown activation");
  }
#undef P2P_BWD_REDUCE
#undef P2P_BWD_REDUCE_U
  P2P_LAUNCH_CHECK();
}

int p2p_bn_bwd_apply(const void* dz, const void* r, const float* sb, const double* red, void* dr, float* dgamma,
                     float* dbeta, float* dbias, int64_t pixels, int C, int c_real, int dtype, void* stream) {
  P2P_REQUIRE(dz && r && sb && red && dr, "null pointer");
  P2P_REQUIRE(C % 16 == 0 && C <= 512 && 256 % (C / 4) == 0, "C must be 16, 32, 64, 128, 256 or 512");
  const int ppb = 256 / (C / 4);
  P2P_DISPATCH(dtype, bn_bwd_apply_kernel<T><<<red_ew_grid(pixels, ppb), 256, 0, as_stream(stream)>>>(
                          (const T*)dz, (const T*)r, sb, red, (T*)dr, dbias, dgamma, dbeta, pixels, C, c_real););
  P2P_LAUNCH_CHECK();
}

int p2p_prelu_grad_finish(const double* dprelu, float* grad, void* stream) {
  P2P_REQUIRE(dprelu && grad, "null pointer");
  prelu_grad_finish_kernel<<<1, 1, 0, as_stream(stream)>>>(dprelu, grad);
  P2P_LAUNCH_CHECK();
}

int p2p_d_input_prep(const float* a, int ca, const float* b, int cb, float* dst, int B, int H, int W, void* stream) {
  P2P_REQUIRE(a && dst && ca > 0 && ca + (b ? cb : 0) <= 8, "at most 8 input channels");
  d_input_prep_kernel<<<ew_grid((int64_t)B * H * W), 256, 0, as_stream(stream)>>>(a, ca, b, b ? cb : 0, dst, B, H, W);
  P2P_LAUNCH_CHECK();
}

int p2p_avgpool3s2_fwd(const float* in, float* out, int B, int H, int W, int C, void* stream) {
  P2P_REQUIRE(in && out && C == 8, "NHWC fp32 with 8 stored channels");
  const int Ho = (H - 1) / 2 + 1, Wo = (W - 1) / 2 + 1;
  avgpool_fwd_kernel<<<ew_grid((int64_t)B * Ho * Wo), 256, 0, as_stream(stream)>>>(in, out, B, H, W, Ho, Wo);
  P2P_LAUNCH_CHECK();
}

int p2p_avgpool3s2_bwd_add(const float* dout, float* din, int B, int H, int W, int C, void* stream) {
  P2P_REQUIRE(dout && din && C == 8, "NHWC fp32 with 8 stored channels");
  const int Ho = (H - 1) / 2 + 1, Wo = (W - 1) / 2 + 1;
  avgpool_bwd_add_kernel<<<ew_grid((int64_t)B * H * W), 256, 0, as_stream(stream)>>>(dout, din, B, H, W, Ho, Wo);
  P2P_LAUNCH_CHECK();
}

int p2p_cast_pad_channels(const float* in, void* out, int64_t pixels, int Cs, int Cd, int dtype, void* stream) {
  P2P_REQUIRE(in && out && Cs % 8 == 0 && Cd % 8 == 0 && Cd >= Cs, "channel counts must be multiples of 8");
  P2P_DISPATCH(dtype, cast_pad_channels_kernel<T><<<ew_grid(pixels * (Cd / 8)), 256, 0, as_stream(stream)>>>(
                          in, (T*)out, pixels, Cs, Cd););
  P2P_LAUNCH_CHECK();
}

int p2p_uncast_add(const void* in, float* out, int64_t pixels, int Cs, int Cd, int dtype, void* stream) {
  P2P_REQUIRE(in && out && Cs % 8 == 0 && Cd % 8 == 0 && Cd >= Cs, "channel counts must be multiples of 8");
  P2P_DISPATCH(dtype, uncast_add_kernel<T><<<ew_grid(pixels * (Cs / 8)), 256, 0, as_stream(stream)>>>(
                          (const T*)in, out, pixels, Cs, Cd););
  P2P_LAUNCH_CHECK();
}

int p2p_d_input_grad(const float* dx, float* db, int ca, int cb, int B, int H, int W, void* stream) {
  P2P_REQUIRE(dx && db && ca + cb <= 8, "at most 8 input channels");
  d_input_grad_kernel<<<ew_grid((int64_t)B * cb * H * W), 256, 0, as_stream(stream)>>>(dx, db, ca, cb, B, H, W);
  P2P_LAUNCH_CHECK();
}

int p2p_act_bwd(const void* out, const void* g_a, const void* g_b, void* dz, float* dbias, int64_t pixels, int C,
                int c_real, int act, int dtype, void* stream) {
  P2P_REQUIRE(dz && (g_a || g_b), "null pointer");
  P2P_REQUIRE(act == P2P_ACT_NONE || ((act == P2P_ACT_LRELU || act == P2P_ACT_RELU) && out),
              "(Leaky)ReLU gradient needs the forward output");
  P2P_REQUIRE(C % 16 == 0 && C <= 512, "C must be a multiple of 16, <= 512");
  const int ppb = 256 / (C / 8);
  P2P_DISPATCH(dtype, act_bwd_kernel<T><<<red_ew_grid(pixels, ppb), 256, 0, as_stream(stream)>>>(
                          (const T*)out, (const T*)g_a, (const T*)g_b, (T*)dz, dbias, pixels, C, c_real, act););
  P2P_LAUNCH_CHECK();
}

int p2p_tap_gather_sum(const void* t, const float* bias, void* out, int N, int Hi, int Wi, int Ho, int Wo, int kh,
                       int kw, int pad, int act, int dtype, void* stream) {
  P2P_REQUIRE(t && out && N > 0 && kh > 0 && kw > 0 && kh * kw <= 16, "bad arguments");
  P2P_REQUIRE(Ho == Hi + 2 * pad - kh + 1 && Wo == Wi + 2 * pad - kw + 1, "stride-1 geometry expected");
  P2P_REQUIRE(act == P2P_ACT_NONE || act == P2P_ACT_LRELU, "epilogue activation must be none or leaky relu");
  const int64_t total = (int64_t)N * Ho * Wo;
  P2P_DISPATCH(dtype, tap_gather_sum_kernel<T><<<ew_grid(total), 256, 0, as_stream(stream)>>>(
                          (const T*)t, bias, (T*)out, N, Hi, Wi, Ho, Wo, kh, kw, pad, act););
  P2P_LAUNCH_CHECK();
}

int p2p_tap_scatter(const void* dy, void* a, int N, int Hi, int Wi, int Ho, int Wo, int kh, int kw, int pad, int dtype,
                    void* stream) {
  P2P_REQUIRE(dy && a && N > 0 && kh > 0 && kw > 0 && kh * kw <= 16, "bad arguments");
  P2P_REQUIRE(Ho == Hi + 2 * pad - kh + 1 && Wo == Wi + 2 * pad - kw + 1, "stride-1 geometry expected");
  const int64_t total = (int64_t)N * Hi * Wi;
  P2P_DISPATCH(dtype, tap_scatter_kernel<T><<<ew_grid(total), 256, 0, as_stream(stream)>>>((const T*)dy, (T*)a, N, Hi, Wi,
                                                                                          Ho, Wo, kh, kw, pad););
  P2P_LAUNCH_CHECK();
}

int p2p_kx_gather_sum(const void* t, const float* bias, void* out, int N, int H, int Wt, int Wo, int kw, int co_real,
                      int Ct, int Co, int dtype, void* stream) {
  P2P_REQUIRE(t && out && N > 0 && H > 0 && kw > 0 && Wt == Wo + kw - 1, "bad geometry");
  P2P_REQUIRE(co_real >= 1 && co_real <= 8 && kw * co_real <= Ct && Ct % 16 == 0 && Co % 16 == 0, "bad channel counts");
  P2P_REQUIRE(H <= 65535 && N <= 65535, "map too large");
  const int smem = dtype == P2P_BF16 ? (256 + kw - 1) * (Ct / 2 + 1) * 4 : 0;
  P2P_REQUIRE(smem <= 48 * 1024, "fat map too wide for the staging buffer");
  const dim3 grid((Wo + 255) / 256, H, N);
  P2P_DISPATCH(dtype, kx_gather_sum_kernel<T><<<grid, 256, smem, as_stream(stream)>>>((const T*)t, bias, (T*)out, H, Wt, Wo,
                                                                                     kw, co_real, Ct, Co););
  P2P_LAUNCH_CHECK();
}

int p2p_kx_scatter(const void* dy, void* a, int N, int H, int Wo, int Wa, int kw, int co_real, int Cs, int Ca, int dtype,
                   void* stream) {
  P2P_REQUIRE(dy && a && N > 0 && H > 0 && kw > 0 && Wa == Wo + kw - 1, "bad geometry");
  P2P_REQUIRE(co_real >= 1 && co_real <= Cs && kw * co_real <= Ca && Cs % 8 == 0 && Ca % 8 == 0, "bad channel counts");
  const int64_t total = (int64_t)N * H * Wa * (Ca / 8);
  P2P_DISPATCH(dtype, kx_scatter_kernel<T><<<ew_grid(total), 256, 0, as_stream(stream)>>>((const T*)dy, (T*)a, N, H, Wo, Wa,
                                                                                         kw, co_real, Cs, Ca););
  P2P_LAUNCH_CHECK();
}

int p2p_c_output(const void* res, const float* x, float* y, float* compressed, int B, int H, int W, int bits, int dtype,
                 void* stream) {
  P2P_REQUIRE(res && x && (y || compressed), "null pointer");
  P2P_REQUIRE(B > 0 && H > 0 && W > 0 && H % 2 == 0 && W % 2 == 0 && bits >= 1 && bits <= 16, "bad geometry");
  const int64_t total = (int64_t)B * (H / 2) * (W / 2);
  P2P_DISPATCH(dtype, c_output_kernel<T><<<ew_grid(total), 256, 0, as_stream(stream)>>>(
                          (const T*)res, x, y, compressed, B, H, W, (float)((1 << bits) - 1)););
  P2P_LAUNCH_CHECK();
}

int p2p_nhwc_to_nchw(const void* in, float* out, int B, int H, int W, int C, int c_real, int dtype, void* stream) {
  P2P_REQUIRE(in && out && c_real <= C, "null pointer");
  P2P_DISPATCH(dtype, nhwc_to_nchw_kernel<T><<<ew_grid((int64_t)B * c_real * H * W), 256, 0, as_stream(stream)>>>(
                          (const T*)in, out, B, H, W, C, c_real););
  P2P_LAUNCH_CHECK();
}

int p2p_nchw_to_nhwc(const float* in, void* out, int B, int H, int W, int C, int c_real, int dtype, void* stream) {
  P2P_REQUIRE(in && out && c_real <= C && C % 8 == 0, "null pointer");
  P2P_DISPATCH(dtype, nchw_to_nhwc_kernel<T><<<ew_grid((int64_t)B * H * W * (C / 8)), 256, 0, as_stream(stream)>>>(
                          in, (T*)out, B, H, W, C, c_real););
  P2P_LAUNCH_CHECK();
}

int p2p_zero_fill(void* p, int64_t bytes, void* stream) {
  P2P_REQUIRE(p != nullptr && bytes > 0 && ((uintptr_t)p % 16 == 0), "16-byte aligned buffer expected");
  const int64_t n16 = bytes / 16;
  int64_t blocks = (n16 + 255) / 256;
  if (blocks < 1) blocks = 1;
  if (blocks > kSMs * 8) blocks = kSMs * 8;
  zero_fill_kernel<<<(int)blocks, 256, 0, as_stream(stream)>>>((uint4*)p, n16, (uint8_t*)p + n16 * 16, (int)(bytes - n16 * 16));
  P2P_LAUNCH_CHECK();
}

int p2p_decode_u8(const uint8_t* src, float* dst, float* quant, int64_t n, int bits, void* stream) {
  P2P_REQUIRE(src && dst && n > 0 && bits >= 1 && bits <= 16, "bad arguments");
  P2P_REQUIRE(((uintptr_t)src % 16 == 0) && ((uintptr_t)dst % 16 == 0) && ((uintptr_t)quant % 16 == 0),
              "buffers must be 16-byte aligned");
  int64_t blocks = (n / 16 + 255) / 256;
  if (blocks < 1) blocks = 1;
  if (blocks > kSMs * 8) blocks = kSMs * 8;
  decode_u8_kernel<<<(int)blocks, 256, 0, as_stream(stream)>>>(src, dst, quant, n, (float)((1 << bits) - 1));
  P2P_LAUNCH_CHECK();
}

int p2p_maxpool2_fwd(const void* in, void* out, int N, int H, int W, int C, int dtype, void* stream) {
  P2P_REQUIRE(in && out && N > 0 && H >= 2 && W >= 2 && H % 2 == 0 && W % 2 == 0 && C % 8 == 0, "bad arguments");
  const int64_t total = (int64_t)N * (H / 2) * (W / 2) * (C / 8);
  P2P_DISPATCH(dtype, maxpool2_fwd_kernel<T><<<ew_grid(total), 256, 0, as_stream(stream)>>>((const T*)in, (T*)out, N, H, W, C););
  P2P_LAUNCH_CHECK();
}

int p2p_maxpool2_bwd(const void* in, const void* dout, void* din, int N, int H, int W, int C, int dtype, void* stream) {
  P2P_REQUIRE(in && dout && din && N > 0 && H >= 2 && W >= 2 && H % 2 == 0 && W % 2 == 0 && C % 8 == 0, "bad arguments");
  const int64_t total = (int64_t)N * (H / 2) * (W / 2) * (C / 8);
  P2P_DISPATCH(dtype, maxpool2_bwd_kernel<T><<<ew_grid(total), 256, 0, as_stream(stream)>>>((const T*)in, (const T*)dout,
                                                                                          (T*)din, N, H, W, C););
  P2P_LAUNCH_CHECK();
}

}  // extern "C"
