// Weight gradient of convolutions with ONE real output channel (the last layer of every discriminator scale:
// 512 -> 1, 4x4, networks.py:783).  As an implicit GEMM it is 1/16 of an MMA tile wide: the tcgen05 kernel spends
// 77 us (B=16, 35x35 map) on 0.16 GFLOP.  It is one memory-bound sweep over the 19 MB input map,
//   dw[tap][ci] += sum_q x[q][ci] * dy[q - tap]        (per-block partials, fp32 atomics),
// done here on the CUDA cores in 52 us.  (The forward and data-gradient of the same layer were tried the same way
// and lost to the tcgen05 kernels -- 82 vs 46 us and 57 vs 34 us -- so they stay there.)
// Same stored layouts as every other conv kernel (NHWC, 16 stored output channels of which 15 are zero).
#include "common.cuh"

namespace p2p {

__device__ __forceinline__ void ldf8(const float* p, float v[8]) { ld8(p, v); }
__device__ __forceinline__ void ldf8(const __nv_bfloat16* p, float v[8]) { ld8(p, v); }

// ---------------------------------------------------------------------------------------------------
// weight gradient: block (x, y) = pixel stripe x, tap group y (4 taps); thread = (8 channels, pixel lane)
// ---------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) conv_co1_wgrad_kernel(const T* __restrict__ x, const T* __restrict__ dy,
                                                             float* __restrict__ acc, int N, int Hi, int Wi, int Ci,
                                                             int Ho, int Wo, int kh, int kw, int pad) {
  const int C8 = Ci >> 3;                            // 256 % C8 == 0, checked by the caller
  const int c8 = threadIdx.x % C8, lane_p = threadIdx.x / C8, ppb = 256 / C8;
  const int taps = kh * kw, t0 = blockIdx.y * 4;
  float a[4][8];
#pragma unroll
  for (int t = 0; t < 4; ++t)
#pragma unroll
    for (int k = 0; k < 8; ++k) a[t][k] = 0.f;
  const int pixels = N * Hi * Wi;
  for (int q = blockIdx.x * ppb + lane_p; q < pixels; q += gridDim.x * ppb) {
    const int ix = q % Wi, iy = (q / Wi) % Hi, n = q / (Wi * Hi);
    float d[4];
    bool any = false;
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const int tap = t0 + t, ky = tap / kw, kx = tap - ky * kw;
      const int oy = iy + pad - ky, ox = ix + pad - kx;
      const bool ok = tap < taps && oy >= 0 && oy < Ho && ox >= 0 && ox < Wo;
      d[t] = ok ? ldf(dy + (((size_t)n * Ho + oy) * Wo + ox) * 16) : 0.f;
      any |= ok;
    }
    if (!any) continue;
    float v[8];
    ldf8(x + (size_t)q * Ci + c8 * 8, v);
#pragma unroll
    for (int t = 0; t < 4; ++t)
#pragma unroll
      for (int k = 0; k < 8; ++k) a[t][k] = fmaf(v[k], d[t], a[t][k]);
  }
  // combine the pixel lanes of the block, then one atomic per (tap, channel): acc is [tap][Ci][16], co = 0
  __shared__ float s_a[256 * 33];
#pragma unroll
  for (int t = 0; t < 4; ++t)
#pragma unroll
    for (int k = 0; k < 8; ++k) s_a[threadIdx.x * 33 + t * 8 + k] = a[t][k];
  __syncthreads();
  for (int e = threadIdx.x; e < 4 * Ci; e += 256) {
    const int t = e / Ci, ci = e - t * Ci;
    if (t0 + t >= taps) continue;
    float s = 0.f;
    for (int l = 0; l < ppb; ++l) s += s_a[(l * C8 + (ci >> 3)) * 33 + t * 8 + (ci & 7)];
    if (s != 0.f) atomicAdd(acc + ((size_t)(t0 + t) * Ci + ci) * 16, s);
  }
}

// ---------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------
bool conv_co1_supported(const p2p_conv* c) {
  const int C8 = c->Ci / 8;
  return c->co_real == 1 && c->Co == 16 && c->stride == 1 && c->kh == 4 && c->kw == 4 && c->Ci % 64 == 0 &&
         c->Ci <= 1024 && 256 % C8 == 0;
}

int conv_wgrad_co1(const p2p_conv* c, const void* in, const void* dout, float* acc, int dtype, cudaStream_t st) {
  const int ppb = 256 / (c->Ci / 8);
  int64_t stripes = ((int64_t)c->N * c->Hi * c->Wi + ppb - 1) / ppb;
  if (stripes > kSMs) stripes = kSMs;
  const dim3 grid((int)stripes, (c->kh * c->kw + 3) / 4);
  trace_kernel("conv_co1_wgrad");
  P2P_DISPATCH(dtype, conv_co1_wgrad_kernel<T><<<grid, 256, 0, st>>>((const T*)in, (const T*)dout, acc, c->N, c->Hi, c->Wi,
                                                                    c->Ci, c->Ho, c->Wo, c->kh, c->kw, c->pad););
  P2P_LAUNCH_CHECK();
}

}  // namespace p2p
