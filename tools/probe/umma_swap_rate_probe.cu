// Hardware probe (not product code): the MMA of the operand-swapped halo kernel -- A = a canonical [128 x 64] weight
// tile, B = N = 256 pixel rows read through a shifted halo descriptor (8-row groups `sbo_rows` halo rows apart, start
// shifted by the filter tap) -- issued back to back by one thread, optionally while `ld_warps` x 4 other warps read
// the OTHER accumulator set out of TMEM with tcgen05.ld (the epilogue's access pattern), and optionally while TMA-like
// traffic is absent.  Answers: does the strided / shifted B operand cost cycles?  do concurrent TMEM reads?
#include "../../p2p-pytorch_b200/csrc/tc_common.cuh"

using namespace p2p;
using namespace p2p::tc;

__global__ void __launch_bounds__(32 * 17) swap_rate_kernel(long long* out, int iters, int sbo_rows, int shifts, int ld_groups,
                                                            int n_cols) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t done;
  __shared__ uint32_t tmem_s;
  __shared__ volatile int stop;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int total = 16384 + 48 * 1024;   // weight tile + halo (34 x 10 x 128 B = 43.5 KB)
  for (int i = threadIdx.x; i < total / 16; i += blockDim.x) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
  if (threadIdx.x == 0) {
    mbar_init(&done, 1);
    stop = 0;
    fence_barrier_init();
  }
  fence_proxy_async();
  if (warp == 0) tmem_alloc(&tmem_s, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_s;
  if (warp == 0) {
    if (lane == 0) {
      const uint32_t idesc = umma_idesc_bf16(128, n_cols, 0, 0);
      const uint32_t w_hi = desc_hi(8 * 128, 128);
      const uint32_t x_hi = desc_hi((uint32_t)(sbo_rows * 128), 128);
      const uint32_t w_lo = desc_lo(smem_u32(smem), 16), x_lo0 = desc_lo(smem_u32(smem + 16384), 16);
      const long long t0 = clock64();
      for (int i = 0; i < iters; ++i) {
        const uint32_t k = (uint32_t)(i & 3);
        const int tap = (i >> 2) % 9;
        const uint32_t sh = shifts ? (uint32_t)(((tap / 3) * sbo_rows + tap % 3) * 128) >> 4 : 0u;
        umma_bf16_split(tmem, w_lo + 2 * k, w_hi, x_lo0 + sh + 2 * k, x_hi, idesc, 1u);
      }
      umma_commit(&done);
      mbar_wait(&done, 0);
      const long long t1 = clock64();
      stop = 1;
      if (blockIdx.x == 0) out[0] = t1 - t0;
    }
    __syncwarp();
  } else if (warp <= 4 * ld_groups) {
    // epilogue-like readers of accumulator set 1 (columns 256..511)
    const int q = warp & 3;
    uint32_t acc = 0;
    long long n = 0;
    while (!stop) {
      for (int j0 = 0; j0 < 256; j0 += 32) {
        uint32_t v[32];
        tmem_ld32(tmem + ((uint32_t)(q * 32) << 16) + 256u + (uint32_t)j0, v);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j) acc ^= v[j];
      }
      ++n;
    }
    if (acc == 0x12345678u) out[7] = n;       // keep the loads alive
    if (blockIdx.x == 0 && warp == 1 && lane == 0) out[1] = n;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}

extern "C" int swap_rate_run(long long* out_dev, int iters, int sbo_rows, int shifts, int ld_groups, int n_cols, int grid) {
  const int smem = 16384 + 48 * 1024 + 2048;
  cudaFuncSetAttribute(swap_rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  swap_rate_kernel<<<grid, 32 * 17, smem>>>(out_dev, iters, sbo_rows, shifts, ld_groups, n_cols);
  return (int)cudaDeviceSynchronize();
}
