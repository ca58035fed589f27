// Hardware probe, part 2 (not product code): the same absolute-address question for 32- and 64-byte swizzle
// rows, and MN-major operands whose "atoms" OVERLAP (LBO = one row), i.e. consecutive convolution taps read
// from one halo tile.
// mode 2: K-major A with RB-byte rows (RB = 32 or 64, matching swizzle): smem holds ROWS x (RB/2) bf16;
//         A' row i = smem row (shift + (i/8)*grp + i%8); K = RB/2; B K-major [16][RB/2].
// mode 3: MN-major A: smem rows (K index) hold RB/2 channels; the 128 M lanes are 128/(RB/2) taps x RB/2
//         channels, tap t reading rows shifted by t (LBO = one row):
//         A'[t*(RB/2)+c][k] = smem[shift + k + t][c]; K = 64 rows (4 MMAs); B MN-major [64][16].
#include "../../p2p-pytorch_b200/csrc/tc_common.cuh"

using namespace p2p;
using namespace p2p::tc;

constexpr int ROWS2 = 192;

__global__ void __launch_bounds__(128) probe2_kernel(const __grid_constant__ CUtensorMap map_a,
                                                     const __grid_constant__ CUtensorMap map_b, float* out, int mode,
                                                     int shift, int grp_rows, int rb) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar, done;
  __shared__ uint32_t tmem_s;
  uint8_t* sA = smem;
  uint8_t* sB = smem + ROWS2 * 128;
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
    const uint32_t bytes = mode == 2 ? ROWS2 * rb + 16 * rb : ROWS2 * rb + 64 * 32;
    mbar_arrive_expect_tx(&bar, bytes);
    tma_load_2d(sA, &map_a, &bar, 0, 0);
    tma_load_2d(sB, &map_b, &bar, 0, 0);
    mbar_wait(&bar, 0);
    tc_fence_after();
    if (mode == 2) {
      const uint32_t idesc = umma_idesc_bf16(128, 16, 0, 0);
      for (int k = 0; k < rb / 32; ++k) {
        uint64_t ad = smem_desc_kmajor(smem_u32(sA) + shift * rb + k * 32, rb);
        ad &= ~(uint64_t(0x3FFF) << 32);
        ad |= (uint64_t)(((uint32_t)(grp_rows * rb) >> 4) & 0x3FFF) << 32;
        const uint64_t bd = smem_desc_kmajor(smem_u32(sB) + k * 32, rb);
        umma_bf16(tmem, ad, bd, idesc, k != 0);
      }
    } else {
      const uint32_t idesc = umma_idesc_bf16(128, 16, 1, 1);
      for (int k = 0; k < 4; ++k) {
        const uint64_t ad = smem_desc_mnmajor(smem_u32(sA) + (shift + k * 16) * rb, rb, rb);  // LBO = one row
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

// a: [ROWS2][rb/2] bf16; b: mode 2 [16][rb/2], mode 3 [64][16]; out [128][16]
extern "C" int probe2_run(const void* a, const void* b, float* out, int mode, int shift, int grp_rows, int rb,
                          void* stream) {
  CUtensorMap ma, mb;
  const uint32_t es[2] = {1, 1};
  {
    const uint64_t dims[2] = {(uint64_t)(rb / 2), (uint64_t)ROWS2};
    const uint64_t str[1] = {(uint64_t)rb};
    const uint32_t box[2] = {(uint32_t)(rb / 2), (uint32_t)ROWS2};
    if (make_map_bf16(&ma, a, 2, dims, str, box, es, rb)) return -1;
  }
  if (mode == 2) {
    const uint64_t dims[2] = {(uint64_t)(rb / 2), 16};
    const uint64_t str[1] = {(uint64_t)rb};
    const uint32_t box[2] = {(uint32_t)(rb / 2), 16};
    if (make_map_bf16(&mb, b, 2, dims, str, box, es, rb)) return -1;
  } else {
    const uint64_t dims[2] = {16, 64};
    const uint64_t str[1] = {32};
    const uint32_t box[2] = {16, 64};
    if (make_map_bf16(&mb, b, 2, dims, str, box, es, 32)) return -1;
  }
  const int smem = ROWS2 * 128 + 4096 + 1024;
  cudaFuncSetAttribute(probe2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  probe2_kernel<<<1, 128, smem, (cudaStream_t)stream>>>(ma, mb, out, mode, shift, grp_rows, rb);
  return (int)cudaGetLastError();
}
