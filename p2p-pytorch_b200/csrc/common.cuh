// Shared helpers for the libp2p_b200 kernels (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdio>
#include <string>

#include "../../include/p2p_b200.h"

namespace p2p {

void set_error(const std::string& s);
void count_launches(int n);   // kernels enqueued so far (reported by bench.py as gpu_launches)
// Which kernel instantiation served a call (p2p_trace_begin / p2p_trace_end): lets a test prove that the shapes it
// checks run on the same instantiations as the benchmark.  A no-op unless a trace is open.
void trace_kernel(const char* fmt, ...);
bool trace_open();

// cudaFuncAttributeMaxDynamicSharedMemorySize is a per-device attribute: remember the largest opt-in made for a
// kernel on each device ordinal (a process may drive several GPUs through the C ABI).  Setting it twice is harmless,
// so concurrent first calls need no lock.
constexpr int kMaxDevices = 64;
struct SmemOptIn {
  int bytes[kMaxDevices] = {};
};
template <typename K>
static inline cudaError_t ensure_dyn_smem(K kernel, SmemOptIn& cache, int smem) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= kMaxDevices || smem > cache.bytes[dev]) {
    e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e == cudaSuccess && dev >= 0 && dev < kMaxDevices) cache.bytes[dev] = smem;
  }
  return e;
}

#define P2P_REQUIRE(cond, msg)                                            \
  do {                                                                    \
    if (!(cond)) {                                                        \
      ::p2p::set_error(std::string(__func__) + ": " + (msg));             \
      return -1;                                                          \
    }                                                                     \
  } while (0)

#define P2P_LAUNCH_CHECK()                                                               \
  do {                                                                                   \
    cudaError_t e_ = cudaGetLastError();                                                 \
    if (e_ != cudaSuccess) {                                                             \
      ::p2p::set_error(std::string(__func__) + ": " + cudaGetErrorString(e_));           \
      return (int)e_;                                                                    \
    }                                                                                    \
    ::p2p::count_launches(1);                                                            \
    return 0;                                                                            \
  } while (0)

#define P2P_CUDA(call)                                                                   \
  do {                                                                                   \
    cudaError_t e_ = (call);                                                             \
    if (e_ != cudaSuccess) {                                                             \
      ::p2p::set_error(std::string(__func__) + ": " + cudaGetErrorString(e_));           \
      return (int)e_;                                                                    \
    }                                                                                    \
  } while (0)

// dtype dispatch: body sees `T`
#define P2P_DISPATCH(dtype, ...)                          \
  do {                                                    \
    if ((dtype) == P2P_F32) {                             \
      using T = float;                                    \
      __VA_ARGS__                                         \
    } else if ((dtype) == P2P_BF16) {                     \
      using T = __nv_bfloat16;                            \
      __VA_ARGS__                                         \
    } else {                                              \
      ::p2p::set_error("unknown dtype");                  \
      return -1;                                          \
    }                                                     \
  } while (0)

static inline cudaStream_t as_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }
static inline int cdiv(int64_t a, int64_t b) { return (int)((a + b - 1) / b); }
static constexpr int kSMs = 148;

// ---- element access ---------------------------------------------------------------------------
__device__ __forceinline__ float ldf(const float* p) { return *p; }
__device__ __forceinline__ float ldf(const __nv_bfloat16* p) { return __bfloat162float(*p); }
__device__ __forceinline__ void stf(float* p, float v) { *p = v; }
__device__ __forceinline__ void stf(__nv_bfloat16* p, float v) { *p = __float2bfloat16_rn(v); }

// 8 consecutive channels <-> float[8]; pointers are 8-element aligned (C % 16 == 0 everywhere)
__device__ __forceinline__ void ld8(const float* p, float v[8]) {
  const float4 a = reinterpret_cast<const float4*>(p)[0];
  const float4 b = reinterpret_cast<const float4*>(p)[1];
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}
__device__ __forceinline__ void ld8(const __nv_bfloat16* p, float v[8]) {
  const uint4 u = *reinterpret_cast<const uint4*>(p);
  const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    v[2 * i] = __uint_as_float(w[i] << 16);
    v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
  }
}
__device__ __forceinline__ void st8(float* p, const float v[8]) {
  reinterpret_cast<float4*>(p)[0] = make_float4(v[0], v[1], v[2], v[3]);
  reinterpret_cast<float4*>(p)[1] = make_float4(v[4], v[5], v[6], v[7]);
}
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&t);
}
__device__ __forceinline__ void st8(__nv_bfloat16* p, const float v[8]) {
  uint4 u;
  u.x = pack_bf16x2(v[0], v[1]); u.y = pack_bf16x2(v[2], v[3]);
  u.z = pack_bf16x2(v[4], v[5]); u.w = pack_bf16x2(v[6], v[7]);
  *reinterpret_cast<uint4*>(p) = u;
}
// value as the stored type would hold it (keeps BN statistics consistent with what is re-read)
__device__ __forceinline__ float round_as(float v, const float*) { return v; }
__device__ __forceinline__ float round_as(float v, const __nv_bfloat16*) { return __bfloat162float(__float2bfloat16_rn(v)); }

// ---- reductions -------------------------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ unsigned long long warp_sum(unsigned long long v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
// sum over the whole block (blockDim.x multiple of 32, <= 1024); result valid in thread 0
template <typename F>
__device__ __forceinline__ F block_sum(F v) {
  __shared__ F red_[32];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  v = warp_sum(v);
  __syncthreads();
  if (lane == 0) red_[w] = v;
  __syncthreads();
  const int nw = (blockDim.x + 31) >> 5;
  v = (threadIdx.x < nw) ? red_[threadIdx.x] : F(0);
  if (w == 0) v = warp_sum(v);
  return v;
}

__device__ __forceinline__ int reflect_idx(int i, int n) {
  // ReflectionPad2d index map: -1 -> 1, n -> n-2 (single reflection, pad < n)
  if (i < 0) i = -i;
  if (i >= n) i = 2 * (n - 1) - i;
  return i;
}

}  // namespace p2p
