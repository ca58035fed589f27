// Hardware probe (not product code): does tcgen05.mma apply the shared-memory swizzle to ABSOLUTE address bits?
// If yes, a TMA-written halo tile can be read through descriptors whose start address is shifted by whole rows
// (any multiple of the row pitch) and whose 8-row groups are SBO != 8*pitch apart -- which is what an
// implicit-GEMM convolution needs to reuse one halo tile for all kh*kw taps.
//
//   mode 0: K-major A.  smem holds ROWS x 64 bf16 (128 B rows, SWIZZLE_128B, written by TMA).
//           D[128 x 16] = A'[128 x 64] * B[16 x 64]^T with A' row i = smem row (shift + (i/8)*grp + i%8).
//   mode 1: MN-major A. smem rows are the K index.  D[64.. x 16]: A'[m][k] = smem row (shift + k), column m;
//           K = 16 per MMA, 4 MMAs (64 rows), M = 128 = two 64-column boxes LBO apart.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -shared -Xcompiler -fPIC -o libprobe.so umma_shift_probe.cu
#include "../../p2p-pytorch_b200/csrc/tc_common.cuh"

namespace p2p {
void set_error(const std::string&) {}
void count_launches(int) {}
}  // namespace p2p
using namespace p2p;
using namespace p2p::tc;

constexpr int ROWS = 192;

__global__ void __launch_bounds__(128) probe_kernel(const __grid_constant__ CUtensorMap map_a,
                                                    const __grid_constant__ CUtensorMap map_a2,
                                                    const __grid_constant__ CUtensorMap map_b, float* out, int mode,
                                                    int shift, int grp_rows) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar, done;
  __shared__ uint32_t tmem_s;
  uint8_t* sA = smem;                    // ROWS x 128 B (box 0)
  uint8_t* sA2 = smem + ROWS * 128;      // ROWS x 128 B (box 1: next 64 columns, MN-major mode only)
  uint8_t* sB = smem + 2 * ROWS * 128;   // K-major B: 16 rows x 128 B; MN-major mode: B is [64 k rows][16 n] -> 64 x 32 B
  const int warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    mbar_init(&done, 1);
    fence_barrier_init();
  }
  if (warp == 0) tmem_alloc(&tmem_s, 32);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_s;
  if (threadIdx.x == 0) {
    const uint32_t bytes = mode == 0 ? ROWS * 128 + 16 * 128 : 2 * ROWS * 128 + 64 * 32;
    mbar_arrive_expect_tx(&bar, bytes);
    tma_load_2d(sA, &map_a, &bar, 0, 0);
    if (mode == 1) tma_load_2d(sA2, &map_a2, &bar, 64, 0);
    tma_load_2d(sB, &map_b, &bar, 0, 0);
    mbar_wait(&bar, 0);
    tc_fence_after();
    if (mode == 0) {
      const uint32_t idesc = umma_idesc_bf16(128, 16, 0, 0);
      for (int k = 0; k < 4; ++k) {
        uint64_t ad = smem_desc_kmajor(smem_u32(sA) + shift * 128 + k * 32, 128);
        // override SBO: 8-row groups are grp_rows rows apart
        ad &= ~(uint64_t(0x3FFF) << 32);
        ad |= (uint64_t)(((uint32_t)(grp_rows * 128) >> 4) & 0x3FFF) << 32;
        const uint64_t bd = smem_desc_kmajor(smem_u32(sB) + k * 32, 128);
        umma_bf16(tmem, ad, bd, idesc, k != 0);
      }
    } else {
      const uint32_t idesc = umma_idesc_bf16(128, 16, 1, 1);
      for (int k = 0; k < 4; ++k) {   // 16 K rows per MMA
        const uint64_t ad = smem_desc_mnmajor(smem_u32(sA) + (shift + k * 16) * 128, 128, ROWS * 128);
        const uint64_t bd = smem_desc_mnmajor(smem_u32(sB) + k * 16 * 32, 32, 64 * 32);
        umma_bf16(tmem, ad, bd, idesc, k != 0);
      }
    }
    umma_commit(&done);
  }
  __syncthreads();
  mbar_wait(&done, 0);
  tc_fence_after();
  uint32_t v[16];
  tmem_ld16(tmem + ((uint32_t)(warp * 32) << 16), v);
  tmem_ld_wait();
  for (int j = 0; j < 16; ++j) out[threadIdx.x * 16 + j] = __uint_as_float(v[j]);
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 32);
}

// a: [ROWS][128] bf16 (mode 0 uses columns 0..63), b: mode 0 [16][64] bf16, mode 1 [64][16] bf16; out: [128][16] fp32
extern "C" int probe_run(const void* a, const void* b, float* out, int mode, int shift, int grp_rows, void* stream) {
  CUtensorMap ma, ma2, mb;
  const uint32_t es[2] = {1, 1};
  {
    const uint64_t dims[2] = {128, (uint64_t)ROWS};
    const uint64_t str[1] = {256};
    const uint32_t box[2] = {64, (uint32_t)ROWS};
    if (make_map_bf16(&ma, a, 2, dims, str, box, es, 128)) return -1;
    if (make_map_bf16(&ma2, a, 2, dims, str, box, es, 128)) return -1;
  }
  if (mode == 0) {
    const uint64_t dims[2] = {64, 16};
    const uint64_t str[1] = {128};
    const uint32_t box[2] = {64, 16};
    if (make_map_bf16(&mb, b, 2, dims, str, box, es, 128)) return -1;
  } else {
    const uint64_t dims[2] = {16, 64};
    const uint64_t str[1] = {32};
    const uint32_t box[2] = {16, 64};
    if (make_map_bf16(&mb, b, 2, dims, str, box, es, 32)) return -1;
  }
  const int smem = 2 * ROWS * 128 + 4096 + 1024;
  cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  probe_kernel<<<1, 128, smem, (cudaStream_t)stream>>>(ma, ma2, mb, out, mode, shift, grp_rows);
  return (int)cudaGetLastError();
}
