// Hardware probe (not product code): how many cycles does one tcgen05.mma (M=128, N, K=16, bf16, operands in
// shared memory) cost when a single thread issues ITERS of them back to back?  One CTA per SM.
#include "../../p2p-pytorch_b200/csrc/tc_common.cuh"

using namespace p2p;
using namespace p2p::tc;

__global__ void __launch_bounds__(128) rate_kernel(long long* out, int n_cols, int iters, int n_acc, int mode, int rb,
                                                   int sbo_rows) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t done;
  __shared__ uint32_t tmem_s;
  const int warp = threadIdx.x >> 5;
  // operands: A 128 rows x 128 B, B 256 rows x 128 B (contents irrelevant: zero them so no NaN slow paths)
  for (int i = threadIdx.x; i < (128 + 256) * 128 / 16; i += 128) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
  if (threadIdx.x == 0) {
    mbar_init(&done, 1);
    fence_barrier_init();
  }
  fence_proxy_async();
  if (warp == 0) tmem_alloc(&tmem_s, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_s;
  if (threadIdx.x == 0) {
    const uint32_t idesc = umma_idesc_bf16(128, n_cols, 0, 0);
    const uint32_t hi = desc_hi((uint32_t)(sbo_rows * rb), rb);      // A: 8-row groups sbo_rows rows apart
    const uint32_t hib = desc_hi((uint32_t)(8 * rb), rb);
    const uint32_t a_lo = desc_lo(smem_u32(smem), 16), b_lo = desc_lo(smem_u32(smem + 128 * 128), 16);
    const long long t0 = clock64();
    if (mode == 0) {
      for (int i = 0; i < iters; ++i) {
        const uint32_t k = rb == 128 ? (uint32_t)(i & 3) : 0u;
        const uint32_t acc = tmem + (uint32_t)((i >> 2) % n_acc) * (uint32_t)n_cols;
        umma_bf16_split(acc, a_lo + 2 * k + (uint32_t)((i & 7) * (rb >> 4)), hi, b_lo + 2 * k, hib, idesc, 1u);
      }
    } else {
      for (int i = 0; i < iters; ++i) {
        const uint32_t k = (uint32_t)(i & 3);
        const uint32_t acc = tmem + (uint32_t)((i >> 2) % n_acc) * (uint32_t)n_cols;
        umma_bf16(acc, smem_desc_kmajor(smem_u32(smem) + k * 32, 128), smem_desc_kmajor(smem_u32(smem + 128 * 128) + k * 32, 128),
                  idesc, 1u);
      }
    }
    const long long t1 = clock64();
    umma_commit(&done);
    mbar_wait(&done, 0);
    const long long t2 = clock64();
    if (blockIdx.x == 0) { out[0] = t1 - t0; out[1] = t2 - t0; }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}

extern "C" int rate_run2(long long* out_dev, int n_cols, int iters, int n_acc, int rb, int sbo_rows, int grid) {
  const int smem = (128 + 256) * 128 + 2048;
  cudaFuncSetAttribute(rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  rate_kernel<<<grid, 128, smem>>>(out_dev, n_cols, iters, n_acc, 0, rb, sbo_rows);
  return (int)cudaDeviceSynchronize();
}

extern "C" int rate_run(long long* out_dev, int n_cols, int iters, int n_acc, int mode, int grid) {
  const int smem = (128 + 256) * 128 + 2048;
  cudaFuncSetAttribute(rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  rate_kernel<<<grid, 128, smem>>>(out_dev, n_cols, iters, n_acc, mode, 128, 8);
  return (int)cudaDeviceSynchronize();
}
