// Parameter-side kernels: weight packing for the two GEMM operand layouts, gradient unpacking into
// PyTorch's [co][ci][kh][kw] parameter layout, and the hand-rolled SpectralNorm of the reference
// (networks.py:525-582): one power iteration per forward, W = W_bar / sigma, sigma differentiable in W_bar.
#include <mutex>

#include "common.cuh"

namespace p2p {

// w [co_real][ci_real][kh][kw] fp32 -> wp [Co][kh][kw][Ci], wt [Ci][kh][kw][Co]   (zeros in the padding)
template <typename T>
__global__ void pack_weights_kernel(const float* __restrict__ w, const float* __restrict__ sigma, T* __restrict__ wp,
                                    T* __restrict__ wt, int co_real, int ci_real, int Co, int Ci, int kh, int kw) {
  const float inv = sigma != nullptr ? 1.f / *sigma : 1.f;
  const int taps = kh * kw;
  const int64_t total = (int64_t)Co * taps * Ci;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    if (wp != nullptr) {  // i enumerates [Co][tap][Ci]
      const int ci = (int)(i % Ci);
      const int tap = (int)((i / Ci) % taps);
      const int co = (int)(i / ((int64_t)Ci * taps));
      const float v = (co < co_real && ci < ci_real) ? w[((int64_t)co * ci_real + ci) * taps + tap] * inv : 0.f;
      stf(wp + i, v);
    }
    if (wt != nullptr) {  // i enumerates [Ci][tap][Co]
      const int co = (int)(i % Co);
      const int tap = (int)((i / Co) % taps);
      const int ci = (int)(i / ((int64_t)Co * taps));
      const float v = (co < co_real && ci < ci_real) ? w[((int64_t)co * ci_real + ci) * taps + tap] * inv : 0.f;
      stf(wt + i, v);
    }
  }
}

// grad[co][ci][tap] += scale * acc[tap][ci][co];  optionally dot += sum acc * w
__global__ void __launch_bounds__(256) unpack_wgrad_kernel(const float* __restrict__ acc, float* __restrict__ grad,
                                                           const float* __restrict__ w, const float* __restrict__ sigma,
                                                           float* __restrict__ dot, int co_real, int ci_real, int Co,
                                                           int Ci, int taps) {
  const float scale = sigma != nullptr ? 1.f / *sigma : 1.f;
  const int64_t total = (int64_t)co_real * ci_real * taps;
  float d = 0.f;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int tap = (int)(i % taps);
    const int ci = (int)((i / taps) % ci_real);
    const int co = (int)(i / ((int64_t)taps * ci_real));
    const float a = acc[((int64_t)tap * Ci + ci) * Co + co];
    grad[i] += a * scale;
    if (dot != nullptr) d += a * w[i];
  }
  if (dot != nullptr) {
    const float tot = block_sum<float>(d);
    if (threadIdx.x == 0) atomicAdd(dot, tot);
  }
}

// d/dW_bar of W = W_bar / sigma, sigma = u^T W_bar v:  gW_bar = gW/sigma - (sum gW*W_bar)/sigma^2 * u v^T.
// `dot` holds sum(gW * W_bar) (gW = gradient w.r.t. the effective weight).
__global__ void sn_weight_grad_kernel(float* __restrict__ grad, const float* __restrict__ dot,
                                      const float* __restrict__ sigma, const float* __restrict__ u,
                                      const float* __restrict__ v, int rows, int cols) {
  const float s = *sigma;
  const float coef = *dot / (s * s);
  const int64_t total = (int64_t)rows * cols;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int k = (int)(i % cols);
    const int r = (int)(i / cols);
    grad[i] -= coef * u[r] * v[k];
  }
}

// ---- batched forms: blockIdx.y = job, the blocks of a row stride over that job's elements -----------------
struct PackTable { p2p_pack_job j[P2P_MULTI_MAX]; };
struct UnpackTable { p2p_unpack_job j[P2P_MULTI_MAX]; };
struct SnGradTable { p2p_sn_grad_job j[P2P_MULTI_MAX]; };

// Both directions are a [columns x taps] <-> [taps x columns] transpose per (row, column chunk), staged in shared
// memory so that global reads AND writes are coalesced runs (a thread-per-element gather is L1-tag bound: measured
// 76 us for the discriminator's 68 MB instead of ~15).  Rows of the tile are padded to taps + 1 floats.
constexpr int kTileFloats = 8192;

__device__ __forceinline__ int tile_cols(int Cc, int taps) {
  int cw = kTileFloats / (taps + 1);
  cw = cw > 256 ? 256 : cw & ~15;
  return Cc < cw ? Cc : cw;
}

// Element e of a [ncol x taps] tile in (column, tap) order and in (tap, column) order, advanced by 256 per step
// without divisions in the loops (a division per element made the first version instruction-bound: IPC 3.1).
struct Walk {
  int a, b, da, db, nb;        // e = a * nb + b;  e += 256  ->  a += da, b += db with carry
  __device__ __forceinline__ Walk(int e0, int nb_) : nb(nb_) { a = e0 / nb_; b = e0 - a * nb_; da = 256 / nb_; db = 256 - da * nb_; }
  __device__ __forceinline__ void step() { a += da; b += db; if (b >= nb) { b -= nb; ++a; } }
};

// dst[(r * taps + t) * Cc + c] = inv * w[co][ci][t]  for the columns c of one chunk;  (co, ci) = (r, c) or (c, r)
template <typename T>
__device__ __forceinline__ void pack_layout(float* tile, const float* __restrict__ w, float inv, T* __restrict__ dst,
                                            int R, int Cc, int r_real, int c_real, int taps, int ci_real,
                                            bool rows_are_co) {
  const int cw = tile_cols(Cc, taps), pitch = taps + 1;
  const int chunks = (Cc + cw - 1) / cw, items = R * chunks;
  const uint32_t cstride = rows_are_co ? (uint32_t)taps : (uint32_t)ci_real * taps;   // floats between columns
  const Walk ld0(threadIdx.x, taps);           // (column, tap)
  const Walk st_full(threadIdx.x, cw);         // (tap, column) for full chunks
  for (int item = blockIdx.x; item < items; item += gridDim.x) {
    const int r = item / chunks, c0 = (item - r * chunks) * cw;
    const int ncol = min(cw, Cc - c0), n = ncol * taps;
    const float* src = rows_are_co ? w + ((uint32_t)r * ci_real + c0) * (uint32_t)taps
                                   : w + ((uint32_t)c0 * ci_real + r) * (uint32_t)taps;
    const bool row_ok = r < r_real;
    __syncthreads();
    Walk ld = ld0;
#pragma unroll 4
    for (int e = threadIdx.x; e < n; e += 256) {
      tile[ld.a * pitch + ld.b] = (row_ok && c0 + ld.a < c_real) ? src[ld.a * cstride + ld.b] * inv : 0.f;
      ld.step();
    }
    __syncthreads();
    T* d = dst + (uint32_t)r * taps * Cc + c0;
    Walk st = ncol == cw ? st_full : Walk(threadIdx.x, ncol);
#pragma unroll 4
    for (int e = threadIdx.x; e < n; e += 256) {
      stf(d + (uint32_t)st.a * Cc + st.b, tile[st.b * pitch + st.a]);
      st.step();
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(256) pack_weights_multi_kernel(const __grid_constant__ PackTable tb) {
  __shared__ float tile[kTileFloats];
  const p2p_pack_job& jb = tb.j[blockIdx.y];
  const float inv = jb.sigma != nullptr ? 1.f / *jb.sigma : 1.f;
  const int taps = jb.kh * jb.kw;
  if (jb.w_packed != nullptr)   // [Co][tap][Ci]
    pack_layout<T>(tile, jb.w, inv, static_cast<T*>(jb.w_packed), jb.Co, jb.Ci, jb.co_real, jb.ci_real, taps, jb.ci_real,
                   true);
  if (jb.wt_packed != nullptr)  // [Ci][tap][Co]
    pack_layout<T>(tile, jb.w, inv, static_cast<T*>(jb.wt_packed), jb.Ci, jb.Co, jb.ci_real, jb.co_real, taps, jb.ci_real,
                   false);
}

// grad[co][ci][:] += scale * acc[:][ci][co]: per (ci, chunk of co) the tile is read along co and written as runs of taps
__global__ void __launch_bounds__(256) unpack_wgrad_multi_kernel(const __grid_constant__ UnpackTable tb) {
  __shared__ float tile[kTileFloats];
  const p2p_unpack_job& jb = tb.j[blockIdx.y];
  const float* __restrict__ acc = jb.dw_acc;
  float* __restrict__ grad = jb.grad;
  const float* __restrict__ w = jb.w;
  const bool want_dot = jb.dot != nullptr;
  const float scale = jb.sigma != nullptr ? 1.f / *jb.sigma : 1.f;
  const int Ci = jb.Ci, Co = jb.Co, taps = jb.kh * jb.kw, ci_real = jb.ci_real, co_real = jb.co_real;
  const int cw = tile_cols(co_real, taps), pitch = taps + 1;
  const int chunks = (co_real + cw - 1) / cw, items = ci_real * chunks;
  const uint32_t tap_stride = (uint32_t)Ci * Co, col_stride = (uint32_t)ci_real * taps;
  const Walk st0(threadIdx.x, taps);           // (column, tap)
  const Walk ld_full(threadIdx.x, cw);         // (tap, column) for full chunks
  float d = 0.f;
  for (int item = blockIdx.x; item < items; item += gridDim.x) {
    const int ci = item / chunks, c0 = (item - ci * chunks) * cw;
    const int ncol = min(cw, co_real - c0), n = ncol * taps;
    const float* a = acc + (uint32_t)ci * Co + c0;
    __syncthreads();
    Walk ld = ncol == cw ? ld_full : Walk(threadIdx.x, ncol);
#pragma unroll 4
    for (int e = threadIdx.x; e < n; e += 256) {
      tile[ld.b * pitch + ld.a] = a[ld.a * tap_stride + ld.b];
      ld.step();
    }
    __syncthreads();
    const uint32_t g0 = ((uint32_t)c0 * ci_real + ci) * (uint32_t)taps;
    Walk st = st0;
#pragma unroll 4
    for (int e = threadIdx.x; e < n; e += 256) {
      const uint32_t g = g0 + st.a * col_stride + st.b;
      const float v = tile[st.a * pitch + st.b];
      grad[g] += v * scale;
      if (want_dot) d += v * w[g];
      st.step();
    }
  }
  if (want_dot) {                                             // uniform per block row
    const float tot = block_sum<float>(d);
    if (threadIdx.x == 0 && tot != 0.f) atomicAdd(jb.dot, tot);
  }
}

__global__ void __launch_bounds__(256) sn_weight_grad_multi_kernel(const __grid_constant__ SnGradTable tb) {
  const p2p_sn_grad_job& jb = tb.j[blockIdx.y];
  const float s = *jb.sigma;
  const float coef = *jb.dot / (s * s);
  const uint32_t cols = jb.cols, total = (uint32_t)jb.rows * cols;
  for (uint32_t i = blockIdx.x * 256 + threadIdx.x; i < total; i += gridDim.x * 256) {
    const uint32_t k = i % cols, r = i / cols;
    jb.grad_wbar[i] -= coef * jb.u[r] * jb.v[k];
  }
}

static inline int multi_grid_x(int64_t max_total) {
  int64_t x = (max_total + 4095) / 4096;                    // one item is a [<=256 columns x taps] tile
  return (int)(x < 1 ? 1 : (x > 592 ? 592 : x));
}

// ---- batched power iteration ---------------------------------------------------------------------
struct SnTable {
  const float* w[P2P_SN_MAX_LAYERS];
  float* u[P2P_SN_MAX_LAYERS];
  float* v[P2P_SN_MAX_LAYERS];
  float* sigma[P2P_SN_MAX_LAYERS];
  int rows[P2P_SN_MAX_LAYERS];
  int cols[P2P_SN_MAX_LAYERS];
  int tmp_off[P2P_SN_MAX_LAYERS];  // offset of this layer's scratch (cols + rows floats)
};

// t[k] += sum over a 32-row slab of W[r][k] u[r]      grid (ceil(cols/256), layers, ceil(max_rows/32));
// t is zeroed first (cudaMemsetAsync); 16x more CTAs than one thread per column over all rows
__global__ void sn_wt_u_kernel(SnTable tb, float* __restrict__ tmp) {
  const int l = blockIdx.y;
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  const int rows = tb.rows[l], cols = tb.cols[l];
  const int r0 = blockIdx.z * 32;
  if (k >= cols || r0 >= rows) return;
  const float* w = tb.w[l] + (int64_t)r0 * cols + k;
  const float* u = tb.u[l] + r0;
  const int nr = min(32, rows - r0);
  float acc = 0.f;
#pragma unroll 8
  for (int r = 0; r < nr; ++r) acc = fmaf(w[(int64_t)r * cols], u[r], acc);
  atomicAdd(tmp + tb.tmp_off[l] + k, acc);
}

// v = t / (|t| + eps)            grid (layers), block 1024
__global__ void __launch_bounds__(1024) sn_norm_v_kernel(SnTable tb, const float* __restrict__ tmp) {
  const int l = blockIdx.x;
  const int cols = tb.cols[l];
  const float* t = tmp + tb.tmp_off[l];
  float ss = 0.f;
  for (int k = threadIdx.x; k < cols; k += blockDim.x) ss += t[k] * t[k];
  __shared__ float nrm;
  const float tot = block_sum<float>(ss);
  if (threadIdx.x == 0) nrm = sqrtf(tot) + 1e-12f;
  __syncthreads();
  float* v = tb.v[l];
  for (int k = threadIdx.x; k < cols; k += blockDim.x) v[k] = t[k] / nrm;
}

// s[r] = sum_k W[r][k] v[k]      one warp per row; grid (ceil(rows/8), layers), block 256
__global__ void __launch_bounds__(256) sn_w_v_kernel(SnTable tb, float* __restrict__ tmp) {
  const int l = blockIdx.y;
  const int rows = tb.rows[l], cols = tb.cols[l];
  const int r = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (r >= rows) return;
  const float* w = tb.w[l] + (int64_t)r * cols;
  const float* v = tb.v[l];
  float acc = 0.f;
  for (int k = lane * 4; k < cols; k += 128) {
    const float4 a = *reinterpret_cast<const float4*>(w + k);
    const float4 b = *reinterpret_cast<const float4*>(v + k);
    acc += a.x * b.x + a.y * b.y + a.z * b.z + a.w * b.w;
  }
  acc = warp_sum(acc);
  if (lane == 0) tmp[tb.tmp_off[l] + cols + r] = acc;
}

// u = s / (|s| + eps); sigma = u . s        grid (layers), block 512
__global__ void __launch_bounds__(512) sn_norm_u_kernel(SnTable tb, const float* __restrict__ tmp) {
  const int l = blockIdx.x;
  const int rows = tb.rows[l], cols = tb.cols[l];
  const float* s = tmp + tb.tmp_off[l] + cols;
  float ss = 0.f;
  for (int r = threadIdx.x; r < rows; r += blockDim.x) ss += s[r] * s[r];
  __shared__ float nrm, tot_s;
  const float tot = block_sum<float>(ss);
  if (threadIdx.x == 0) { tot_s = tot; nrm = sqrtf(tot) + 1e-12f; }
  __syncthreads();
  float* u = tb.u[l];
  for (int r = threadIdx.x; r < rows; r += blockDim.x) u[r] = s[r] / nrm;
  if (threadIdx.x == 0) *tb.sigma[l] = tot_s / nrm;
}

}  // namespace p2p

using namespace p2p;

extern "C" {

int p2p_pack_weights(const float* w, const float* sigma, void* w_packed, void* wt_packed, int co_real, int ci_real,
                     int Co, int Ci, int kh, int kw, int dtype, void* stream) {
  P2P_REQUIRE(w && (w_packed || wt_packed), "null pointer");
  P2P_REQUIRE(co_real <= Co && ci_real <= Ci && Co % 16 == 0 && Ci % 16 == 0, "stored channels must be multiples of 16");
  const int64_t total = (int64_t)Co * kh * kw * Ci;
  int64_t blocks = (total + 255) / 256;
  if (blocks > kSMs * 16) blocks = kSMs * 16;
  P2P_DISPATCH(dtype, pack_weights_kernel<T><<<(int)blocks, 256, 0, as_stream(stream)>>>(
                          w, sigma, (T*)w_packed, (T*)wt_packed, co_real, ci_real, Co, Ci, kh, kw););
  P2P_LAUNCH_CHECK();
}

int p2p_unpack_wgrad(const float* dw_acc, float* grad, const float* w, const float* sigma, float* dot, int co_real,
                     int ci_real, int Co, int Ci, int kh, int kw, void* stream) {
  P2P_REQUIRE(dw_acc && grad, "null pointer");
  P2P_REQUIRE(dot == nullptr || w != nullptr, "dot needs the weights");
  const int64_t total = (int64_t)co_real * ci_real * kh * kw;
  int64_t blocks = (total + 255) / 256;
  if (blocks > kSMs * 8) blocks = kSMs * 8;
  unpack_wgrad_kernel<<<(int)blocks, 256, 0, as_stream(stream)>>>(dw_acc, grad, w, sigma, dot, co_real, ci_real, Co, Ci,
                                                                 kh * kw);
  P2P_LAUNCH_CHECK();
}

int p2p_sn_weight_grad(float* grad_wbar, const float* dot, const float* sigma, const float* u, const float* v, int rows,
                       int cols, void* stream) {
  P2P_REQUIRE(grad_wbar && dot && sigma && u && v, "null pointer");
  const int64_t total = (int64_t)rows * cols;
  int64_t blocks = (total + 255) / 256;
  if (blocks > kSMs * 8) blocks = kSMs * 8;
  sn_weight_grad_kernel<<<(int)blocks, 256, 0, as_stream(stream)>>>(grad_wbar, dot, sigma, u, v, rows, cols);
  P2P_LAUNCH_CHECK();
}

int p2p_pack_weights_multi(int n, const p2p_pack_job* jobs, int dtype, void* stream) {
  P2P_REQUIRE(jobs && n >= 1 && n <= P2P_MULTI_MAX, "1..24 jobs per call");
  PackTable tb;
  int64_t mx = 0;
  for (int i = 0; i < n; ++i) {
    const p2p_pack_job& j = jobs[i];
    P2P_REQUIRE(j.w && (j.w_packed || j.wt_packed), "null pointer in a job");
    P2P_REQUIRE(j.co_real <= j.Co && j.ci_real <= j.Ci && j.Co % 16 == 0 && j.Ci % 16 == 0, "bad channel counts");
    const int64_t total = (int64_t)j.Co * j.kh * j.kw * j.Ci;
    P2P_REQUIRE(total < (1LL << 31), "weight tensor too large");
    mx = total > mx ? total : mx;
    tb.j[i] = j;
  }
  P2P_DISPATCH(dtype, pack_weights_multi_kernel<T><<<dim3(multi_grid_x(mx), n), 256, 0, as_stream(stream)>>>(tb););
  P2P_LAUNCH_CHECK();
}

int p2p_unpack_wgrad_multi(int n, const p2p_unpack_job* jobs, void* stream) {
  P2P_REQUIRE(jobs && n >= 1 && n <= P2P_MULTI_MAX, "1..24 jobs per call");
  UnpackTable tb;
  int64_t mx = 0;
  for (int i = 0; i < n; ++i) {
    const p2p_unpack_job& j = jobs[i];
    P2P_REQUIRE(j.dw_acc && j.grad, "null pointer in a job");
    P2P_REQUIRE(j.dot == nullptr || j.w != nullptr, "dot needs the weights");
    const int64_t total = (int64_t)j.Co * j.kh * j.kw * j.Ci;
    P2P_REQUIRE(total < (1LL << 31), "weight tensor too large");
    mx = total > mx ? total : mx;
    tb.j[i] = j;
  }
  unpack_wgrad_multi_kernel<<<dim3(multi_grid_x(mx), n), 256, 0, as_stream(stream)>>>(tb);
  P2P_LAUNCH_CHECK();
}

int p2p_sn_weight_grad_multi(int n, const p2p_sn_grad_job* jobs, void* stream) {
  P2P_REQUIRE(jobs && n >= 1 && n <= P2P_MULTI_MAX, "1..24 jobs per call");
  SnGradTable tb;
  int64_t mx = 0;
  for (int i = 0; i < n; ++i) {
    const p2p_sn_grad_job& j = jobs[i];
    P2P_REQUIRE(j.grad_wbar && j.dot && j.sigma && j.u && j.v && j.rows > 0 && j.cols > 0, "null pointer in a job");
    const int64_t total = (int64_t)j.rows * j.cols;
    P2P_REQUIRE(total < (1LL << 31), "weight tensor too large");
    mx = total > mx ? total : mx;
    tb.j[i] = j;
  }
  sn_weight_grad_multi_kernel<<<dim3(multi_grid_x(mx), n), 256, 0, as_stream(stream)>>>(tb);
  P2P_LAUNCH_CHECK();
}

int p2p_spectral_norm(int n_layers, const float* const* w_host, float* const* u_host, float* const* v_host,
                      float* const* sigma_host, const int32_t* rows_host, const int32_t* cols_host, float* tmp,
                      void* stream) {
  P2P_REQUIRE(n_layers >= 1 && n_layers <= P2P_SN_MAX_LAYERS, "1..16 layers per call");
  P2P_REQUIRE(w_host && u_host && v_host && sigma_host && rows_host && cols_host && tmp, "null pointer");
  SnTable tb;
  int off = 0, max_rows = 0, max_cols = 0;
  for (int l = 0; l < n_layers; ++l) {
    P2P_REQUIRE(w_host[l] && u_host[l] && v_host[l] && sigma_host[l], "null layer pointer");
    P2P_REQUIRE(cols_host[l] % 4 == 0, "cols must be a multiple of 4");
    tb.w[l] = w_host[l]; tb.u[l] = u_host[l]; tb.v[l] = v_host[l]; tb.sigma[l] = sigma_host[l];
    tb.rows[l] = rows_host[l]; tb.cols[l] = cols_host[l]; tb.tmp_off[l] = off;
    off += rows_host[l] + cols_host[l];
    max_rows = rows_host[l] > max_rows ? rows_host[l] : max_rows;
    max_cols = cols_host[l] > max_cols ? cols_host[l] : max_cols;
  }
  cudaStream_t st = as_stream(stream);
  P2P_CUDA(cudaMemsetAsync(tmp, 0, sizeof(float) * (size_t)off, st));
  sn_wt_u_kernel<<<dim3(cdiv(max_cols, 256), n_layers, cdiv(max_rows, 32)), 256, 0, st>>>(tb, tmp);
  sn_norm_v_kernel<<<n_layers, 1024, 0, st>>>(tb, tmp);
  sn_w_v_kernel<<<dim3(cdiv(max_rows, 8), n_layers), 256, 0, st>>>(tb, tmp);
  sn_norm_u_kernel<<<n_layers, 512, 0, st>>>(tb, tmp);
  count_launches(3);
  P2P_LAUNCH_CHECK();
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------
// Allocation-free variant of the power iteration.
//
// p2p_spectral_norm needs sum(rows + cols) floats of scratch that its caller has to provide, i.e. allocate.
// Some callers of this library cannot allocate device memory at the point where they run the
// discriminator: a host program that was handed a fixed memory budget and carved all of its buffers out
// up front (serving loops, embedded trainers), and latency-sensitive loops in which a cudaMalloc/cudaFree
// pair would synchronise the whole device in the middle of a step.  For them the scratch below is part of
// the library image itself: a fixed-size array in device memory that every GPU which loads the library
// holds from load until the process exits.
//
// g_sn_scratch as written in HOST code is only the host-side handle of the symbol -- it is not a device
// address and must never be handed to a kernel.  The device address is looked up once per device with
// cudaGetSymbolAddress (with that device current) and cached.
// ---------------------------------------------------------------------------------------------------
static __device__ __align__(256) float g_sn_scratch[P2P_SN_STATIC_MAX_FLOATS];

namespace p2p {
static float* g_sn_scratch_addr[kMaxDevices] = {};
static std::mutex g_sn_scratch_mutex;

// device address of g_sn_scratch on the current device; the first call per device resolves it
static int sn_static_scratch(float** out) {
  int dev = -1;
  P2P_CUDA(cudaGetDevice(&dev));
  P2P_REQUIRE(dev >= 0 && dev < kMaxDevices, "device ordinal out of range");
  std::lock_guard<std::mutex> lock(g_sn_scratch_mutex);
  if (g_sn_scratch_addr[dev] == nullptr) {
    float* p = nullptr;
    P2P_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_sn_scratch));
    P2P_REQUIRE(p != nullptr, "cudaGetSymbolAddress returned a null address");
    g_sn_scratch_addr[dev] = p;
  }
  *out = g_sn_scratch_addr[dev];
  return 0;
}
}  // namespace p2p

extern "C" int p2p_spectral_norm_static(int n_layers, const float* const* w_host, float* const* u_host,
                                        float* const* v_host, float* const* sigma_host, const int32_t* rows_host,
                                        const int32_t* cols_host, void* stream) {
  P2P_REQUIRE(n_layers >= 1 && n_layers <= P2P_SN_MAX_LAYERS, "1..16 layers per call");
  P2P_REQUIRE(rows_host && cols_host, "null pointer");
  long long need = 0;
  for (int l = 0; l < n_layers; ++l) {
    P2P_REQUIRE(rows_host[l] > 0 && cols_host[l] > 0, "empty weight matrix");
    need += (long long)rows_host[l] + cols_host[l];
  }
  P2P_REQUIRE(need <= P2P_SN_STATIC_MAX_FLOATS,
              "sum(rows + cols) = " + std::to_string(need) + " floats exceeds the built-in scratch of " +
                  std::to_string(P2P_SN_STATIC_MAX_FLOATS) + " floats: use p2p_spectral_norm with your own buffer");
  float* scratch = nullptr;
  if (int e = sn_static_scratch(&scratch)) return e;
  // The array still holds whatever the previous call left in it.  The W^T u partial sums are accumulated with
  // atomics and must start from zero: p2p_spectral_norm clears the part of the scratch it uses (a memset on
  // `stream`, no allocation) before its first kernel; W v then overwrites its slice before anything reads it.
  return p2p_spectral_norm(n_layers, w_host, u_host, v_host, sigma_host, rows_host, cols_host, scratch, stream);
}
