// CUDA-core implicit-GEMM convolutions with fp32 accumulation: the fp32 verification mode and the
// route for shapes the tcgen05 path does not take.  Same tensor layouts as conv_tc.cu.
//   fwd  : M = N*Ho*Wo pixels, Ncol = Co, K = kh*kw*Ci       (reference networks.py:404,422,766-784)
//   dgrad: M = N*Hi*Wi pixels, Ncol = Ci, K = kh*kw*Co
//   wgrad: M = kh*kw*Ci,       Ncol = Co, K = N*Ho*Wo pixels (split over pixel chunks, fp32 atomics)
#include "common.cuh"

namespace p2p {

constexpr int TM = 64, TN = 64, TK = 16;  // block tile; 256 threads, 4x4 outputs each

struct ConvDims {
  int N, Hi, Wi, Ci, Ho, Wo, Co, kh, kw, stride, pad, co_real, act;
};

static ConvDims dims_of(const p2p_conv* c) {
  return ConvDims{c->N, c->Hi, c->Wi, c->Ci, c->Ho, c->Wo, c->Co, c->kh, c->kw, c->stride, c->pad, c->co_real, c->act};
}

template <typename T>
__device__ __forceinline__ void ld4(const T* p, float v[4]);
template <>
__device__ __forceinline__ void ld4<float>(const float* p, float v[4]) {
  const float4 a = *reinterpret_cast<const float4*>(p);
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w;
}
template <>
__device__ __forceinline__ void ld4<__nv_bfloat16>(const __nv_bfloat16* p, float v[4]) {
  const uint2 u = *reinterpret_cast<const uint2*>(p);
  v[0] = __uint_as_float(u.x << 16); v[1] = __uint_as_float(u.x & 0xffff0000u);
  v[2] = __uint_as_float(u.y << 16); v[3] = __uint_as_float(u.y & 0xffff0000u);
}

// DGRAD == false: A row m is output pixel (n,oy,ox), reads in[n, oy*s+ky-pad, ox*s+kx-pad, :Ci]
// DGRAD == true : A row m is input pixel (n,y,x),   reads dout[n, (y+pad-ky)/s, (x+pad-kx)/s, :Co]
template <typename T, bool DGRAD>
__global__ void __launch_bounds__(256) conv_simt_kernel(const T* __restrict__ a, const T* __restrict__ w,
                                                        const float* __restrict__ bias, T* __restrict__ out,
                                                        ConvDims d) {
  __shared__ float As[TK][TM + 4];
  __shared__ float Bs[TK][TN + 4];
  const int Hm = DGRAD ? d.Hi : d.Ho, Wm = DGRAD ? d.Wi : d.Wo;       // pixel grid of the rows
  const int Hs = DGRAD ? d.Ho : d.Hi, Ws = DGRAD ? d.Wo : d.Wi;       // grid of the tensor being read
  const int Ck = DGRAD ? d.Co : d.Ci;                                 // reduction channels
  const int Cn = DGRAD ? d.Ci : d.Co;                                 // output channels
  const int64_t M = (int64_t)d.N * Hm * Wm;
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  const int64_t m0 = (int64_t)blockIdx.x * TM;
  const int n0 = blockIdx.y * TN;

  // loader role: row lr (pixel / out-channel) and 4 consecutive reduction channels lc..lc+3
  const int lr = tid >> 2, lc = (tid & 3) * 4;
  const int64_t pm = m0 + lr;
  const bool pvalid = pm < M;
  int pn = 0, py = 0, px = 0;
  if (pvalid) {
    pn = (int)(pm / ((int64_t)Hm * Wm));
    const int rem = (int)(pm - (int64_t)pn * Hm * Wm);
    py = rem / Wm;
    px = rem - py * Wm;
  }
  const int wrow = n0 + lr;  // weight row (output channel of this GEMM)
  const bool wvalid = wrow < Cn;

  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  for (int ky = 0; ky < d.kh; ++ky) {
    for (int kx = 0; kx < d.kw; ++kx) {
      bool ok = pvalid;
      int sy, sx;
      if (!DGRAD) {
        sy = py * d.stride + ky - d.pad;
        sx = px * d.stride + kx - d.pad;
      } else {
        const int ty_ = py + d.pad - ky, tx_ = px + d.pad - kx;
        ok = ok && ty_ >= 0 && tx_ >= 0 && (ty_ % d.stride) == 0 && (tx_ % d.stride) == 0;
        sy = ty_ / d.stride;
        sx = tx_ / d.stride;
      }
      ok = ok && sy >= 0 && sy < Hs && sx >= 0 && sx < Ws;
      const T* arow = a + (((int64_t)pn * Hs + sy) * Ws + sx) * Ck;
      const T* wrow_p = w + (((int64_t)wrow * d.kh + ky) * d.kw + kx) * Ck;
      for (int c0 = 0; c0 < Ck; c0 += TK) {
        float av[4] = {0.f, 0.f, 0.f, 0.f}, bv[4] = {0.f, 0.f, 0.f, 0.f};
        if (ok) ld4<T>(arow + c0 + lc, av);
        if (wvalid) ld4<T>(wrow_p + c0 + lc, bv);
        __syncthreads();
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          As[lc + i][lr] = av[i];
          Bs[lc + i][lr] = bv[i];
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < TK; ++k) {
          const float4 a4 = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
          const float4 b4 = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
          const float ar[4] = {a4.x, a4.y, a4.z, a4.w}, br[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
          for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(ar[i], br[j], acc[i][j]);
        }
      }
    }
  }
  // epilogue
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int64_t m = m0 + ty * 4 + i;
    if (m >= M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = n0 + tx * 4 + j;
      if (n >= Cn) continue;
      float v = acc[i][j];
      if (!DGRAD) {
        if (bias != nullptr && n < d.co_real) v += bias[n];
        if (d.act == P2P_ACT_LRELU) v = v > 0.f ? v : 0.2f * v;
        if (d.act == P2P_ACT_RELU) v = v > 0.f ? v : 0.f;
        if (n >= d.co_real) v = 0.f;
      }
      stf(out + m * Cn + n, v);
    }
  }
}

// wgrad: rows r = (tap, ci), cols = co, reduction over a chunk of output pixels
template <typename T>
__global__ void __launch_bounds__(256) conv_wgrad_simt_kernel(const T* __restrict__ in, const T* __restrict__ dout,
                                                              float* __restrict__ dw, ConvDims d, int pix_per_chunk) {
  __shared__ float As[TK][TM + 4];
  __shared__ float Bs[TK][TN + 4];
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  const int R = d.kh * d.kw * d.Ci;
  const int r0 = blockIdx.x * TM, n0 = blockIdx.y * TN;
  const int64_t P = (int64_t)d.N * d.Ho * d.Wo;
  const int64_t pbeg = (int64_t)blockIdx.z * pix_per_chunk;
  const int64_t pend = min(P, pbeg + pix_per_chunk);

  // loader: pixel slot lp (0..15), 4 consecutive rows / cols
  const int lp = tid >> 4, l4 = (tid & 15) * 4;
  const int r = r0 + l4;
  const bool rvalid = r < R;
  const int tap = rvalid ? r / d.Ci : 0, ci = rvalid ? r % d.Ci : 0;
  const int ky = tap / d.kw, kx = tap % d.kw;
  const bool cvalid = (n0 + l4) < d.Co;

  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  for (int64_t p0 = pbeg; p0 < pend; p0 += TK) {
    const int64_t p = p0 + lp;
    float av[4] = {0.f, 0.f, 0.f, 0.f}, bv[4] = {0.f, 0.f, 0.f, 0.f};
    if (p < pend) {
      const int n = (int)(p / ((int64_t)d.Ho * d.Wo));
      const int rem = (int)(p - (int64_t)n * d.Ho * d.Wo);
      const int oy = rem / d.Wo, ox = rem - oy * d.Wo;
      const int sy = oy * d.stride + ky - d.pad, sx = ox * d.stride + kx - d.pad;
      if (rvalid && sy >= 0 && sy < d.Hi && sx >= 0 && sx < d.Wi)
        ld4<T>(in + (((int64_t)n * d.Hi + sy) * d.Wi + sx) * d.Ci + ci, av);
      if (cvalid) ld4<T>(dout + p * d.Co + n0 + l4, bv);
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      As[lp][l4 + i] = av[i];
      Bs[lp][l4 + i] = bv[i];
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < TK; ++k) {
      const float4 a4 = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
      const float4 b4 = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
      const float ar[4] = {a4.x, a4.y, a4.z, a4.w}, br[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(ar[i], br[j], acc[i][j]);
    }
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int rr = r0 + ty * 4 + i;
    if (rr >= R) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = n0 + tx * 4 + j;
      if (n < d.Co) atomicAdd(dw + (int64_t)rr * d.Co + n, acc[i][j]);
    }
  }
}

int conv_fwd_simt(const p2p_conv* c, const void* in, const void* w, const float* bias, void* out, int dtype,
                  cudaStream_t st) {
  const ConvDims d = dims_of(c);
  const int64_t M = (int64_t)d.N * d.Ho * d.Wo;
  dim3 grid(cdiv(M, TM), cdiv(d.Co, TN));
  trace_kernel("conv_simt fwd dtype=%d", dtype);
  P2P_DISPATCH(dtype, conv_simt_kernel<T, false><<<grid, 256, 0, st>>>((const T*)in, (const T*)w, bias, (T*)out, d););
  P2P_LAUNCH_CHECK();
}

int conv_dgrad_simt(const p2p_conv* c, const void* dout, const void* wt, void* din, int dtype, cudaStream_t st) {
  const ConvDims d = dims_of(c);
  const int64_t M = (int64_t)d.N * d.Hi * d.Wi;
  dim3 grid(cdiv(M, TM), cdiv(d.Ci, TN));
  trace_kernel("conv_simt dgrad dtype=%d", dtype);
  P2P_DISPATCH(dtype,
               conv_simt_kernel<T, true><<<grid, 256, 0, st>>>((const T*)dout, (const T*)wt, nullptr, (T*)din, d););
  P2P_LAUNCH_CHECK();
}

int conv_wgrad_simt(const p2p_conv* c, const void* in, const void* dout, float* dw, int dtype, cudaStream_t st) {
  const ConvDims d = dims_of(c);
  const int R = d.kh * d.kw * d.Ci;
  const int64_t P = (int64_t)d.N * d.Ho * d.Wo;
  const int gx = cdiv(R, TM), gy = cdiv(d.Co, TN);
  int chunks = (int)((4LL * kSMs + (int64_t)gx * gy - 1) / ((int64_t)gx * gy));  // ~4 CTAs per SM in total
  chunks = (int)min((int64_t)max(chunks, 1), (P + 255) / 256);
  chunks = max(1, min(chunks, 65535));
  int ppc = (int)((P + chunks - 1) / chunks);
  ppc = ((ppc + TK - 1) / TK) * TK;
  chunks = (int)((P + ppc - 1) / ppc);
  dim3 grid(gx, gy, chunks);
  trace_kernel("conv_wgrad_simt dtype=%d", dtype);
  P2P_DISPATCH(dtype, conv_wgrad_simt_kernel<T><<<grid, 256, 0, st>>>((const T*)in, (const T*)dout, dw, d, ppc););
  P2P_LAUNCH_CHECK();
}

}  // namespace p2p
