// sm_100a building blocks for the tensor-core convolutions: mbarrier, TMA (cp.async.bulk.tensor),
// tcgen05 (alloc / mma / commit / ld), UMMA shared-memory and instruction descriptors, and the host-side
// tensor-map encoder (driver entry point fetched at run time, libcuda is not linked).
#pragma once
#include <cuda.h>

#include "common.cuh"

namespace p2p {
namespace tc {

// ---------------------------------------------------------------------------------------------------
// host: tensor maps
// ---------------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn == nullptr) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

inline CUtensorMapSwizzle swizzle_for_bytes(int inner_bytes) {
  return inner_bytes >= 128 ? CU_TENSOR_MAP_SWIZZLE_128B
                            : (inner_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B);
}

// bf16 tensor map of rank 2..4; dims/strides innermost first, strides in BYTES for dims 1..rank-1.
inline int make_map_bf16(CUtensorMap* m, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_b,
                         const uint32_t* box, const uint32_t* estr, int inner_bytes) {
  EncodeTiledFn fn = encode_fn();
  if (fn == nullptr) {
    set_error("cuTensorMapEncodeTiled not available from the driver");
    return -1;
  }
  cuuint64_t gd[5], gs[5];
  cuuint32_t bx[5], es[5];
  for (int i = 0; i < rank; ++i) { gd[i] = dims[i]; bx[i] = box[i]; es[i] = estr[i]; }
  for (int i = 0; i + 1 < rank; ++i) gs[i] = strides_b[i];
  CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base), gd, gs, bx, es,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle_for_bytes(inner_bytes), CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed with CUresult " + std::to_string((int)r));
    return (int)r;
  }
  return 0;
}

// ---------------------------------------------------------------------------------------------------
// device: PTX wrappers
// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ bool elect_one() {
  uint32_t pred = 0;
  asm volatile(
      "{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.u32 %0, 1, 0, P;\n\t}\n" : "=r"(pred));
  return pred != 0;
}

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred P;\n\tmbarrier.try_wait.parity.shared::cta.b64 P, [%1], %2;\n\tselp.u32 %0, 1, 0, P;\n\t}\n"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// Bounded wait: a pipeline bug must surface as a launch failure, never as a hung GPU.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  for (uint32_t spin = 0; !mbar_try_wait(bar, parity); ++spin) {
    if (spin > (1u << 26)) {
      printf("p2p_b200: mbarrier wait timed out (block %d,%d thread %d)\n", blockIdx.x, blockIdx.y, threadIdx.x);
      __trap();
    }
  }
}

__device__ __forceinline__ void prefetch_tmap(const CUtensorMap* m) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(m)) : "memory");
}
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
          smem_u32(dst)),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_load_4d(void* dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2,
                                            int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], "
      "[%2];" ::"r"(smem_u32(dst)),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}

__device__ __forceinline__ void tmem_alloc(uint32_t* dst_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// D[tmem] (+)= A[smem] * B[smem], bf16 inputs, fp32 accumulate; issued by ONE thread
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                          uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive on an mbarrier when all previously issued MMAs of this thread have completed
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}
// 32 lanes x 32 consecutive fp32 columns -> 32 registers per thread (thread = lane = accumulator row)
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t v[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
        "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
        "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t v[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// named barrier over a subset of the CTA's warps (id 1..15; 0 is __syncthreads)
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// ---- descriptors (cute/arch/mma_sm100_desc.hpp bit layout) -------------------------------------------
// UMMA LayoutType codes: 0 none, 2 = 128B swizzle, 4 = 64B, 6 = 32B
__host__ __device__ constexpr uint32_t umma_layout_for_bytes(int inner_bytes) {
  return inner_bytes >= 128 ? 2u : (inner_bytes == 64 ? 4u : 6u);
}
// K-major operand tile whose rows are `row_bytes` (= swizzle span) long and dense: 8-row groups are
// 8*row_bytes apart (SBO); LBO is unused for swizzled K-major layouts (canonical value 1).
__device__ __forceinline__ uint64_t smem_desc_kmajor(uint32_t saddr, int row_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3FFF);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(((uint32_t)(8 * row_bytes) >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;  // descriptor version for sm_100
  d |= (uint64_t)umma_layout_for_bytes(row_bytes) << 61;
  return d;
}
// MN-major operand (the reduction index runs over smem ROWS, the M/N index is contiguous inside a row):
// rows are `row_bytes` (= swizzle span, 64 elements at 128 B) long; 8-row groups are 8*row_bytes apart (SBO);
// the next block of M/N elements (next `row_bytes/2` elements) starts `lbo_bytes` further on (LBO).
__device__ __forceinline__ uint64_t smem_desc_mnmajor(uint32_t saddr, int row_bytes, uint32_t lbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)(((uint32_t)(8 * row_bytes) >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)umma_layout_for_bytes(row_bytes) << 61;
  return d;
}
// kind::f16 instruction descriptor: bf16 x bf16 -> fp32, M x N tile; majors: 0 = K-major, 1 = MN-major
__host__ __device__ constexpr uint32_t umma_idesc_bf16(int M, int N, int a_mn_major, int b_mn_major) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)a_mn_major << 15) | ((uint32_t)b_mn_major << 16) |
         ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

}  // namespace tc
}  // namespace p2p

// ---------------------------------------------------------------------------------------------------
// Shared epilogue of the convolution kernels: one 128 x BLOCK_N fp32 accumulator tile in TMEM ->
// (+bias, LeakyReLU) -> bf16 staging in shared memory -> per-channel sum / sum of squares (BatchNorm
// statistics of the tile as stored) -> coalesced 16-byte global stores.  Called by the 4 epilogue warps
// (128 threads, `et` = 0..127); named barrier 1 is used between the phases.
// ---------------------------------------------------------------------------------------------------
namespace p2p {
namespace tc {

struct EpiOut {
  __nv_bfloat16* out;               // already offset to the image (n * out_sN)
  int64_t out_sH, out_sW;
  int h_mul, h_off, w_mul, w_off;
  int Hv, Wv;                       // valid output-position grid
  int oy0, ox0, tw_log2;            // tile origin; row r of the tile is pixel (oy0 + (r >> tw_log2), ox0 + (r & (tw-1)))
  int n0, co_real, Cn, act;
  double* stats;
};

template <int BLOCK_N>
__device__ __forceinline__ void epilogue_tile(uint32_t tmem_tile, int q, int lane, int et, uint8_t* stg,
                                              const float* s_bias, float* s_red, const EpiOut& e,
                                              uint64_t* release_bar) {
  constexpr int PITCH = BLOCK_N * 2 + 16;
  constexpr int CH = BLOCK_N >= 32 ? 32 : 16;
  const int row = q * 32 + lane;
  const int tw_mask = (1 << e.tw_log2) - 1;
  const bool valid = (e.oy0 + (row >> e.tw_log2)) < e.Hv && (e.ox0 + (row & tw_mask)) < e.Wv;
  const bool full_n = e.n0 + BLOCK_N <= e.co_real;
  const bool lrelu = e.act == P2P_ACT_LRELU || e.act == P2P_ACT_RELU;     // max(f, slope * f)
  const float slope = e.act == P2P_ACT_LRELU ? 0.2f : 0.f;
  // ---- phase 1: TMEM -> registers -> bf16 -> staging ---------------------------------------------
#pragma unroll 1
  for (int c0 = 0; c0 < BLOCK_N; c0 += CH) {
    uint32_t v[32];
    const uint32_t taddr = tmem_tile + ((uint32_t)(q * 32) << 16) + (uint32_t)c0;
    if (CH == 32) tmem_ld32(taddr, v); else tmem_ld16(taddr, v);
    tmem_ld_wait();
    uint32_t pk[CH / 2];
#pragma unroll
    for (int j = 0; j < CH; j += 4) {
      const float4 b4 = *reinterpret_cast<const float4*>(s_bias + c0 + j);
      float f0 = __uint_as_float(v[j]) + b4.x, f1 = __uint_as_float(v[j + 1]) + b4.y;
      float f2 = __uint_as_float(v[j + 2]) + b4.z, f3 = __uint_as_float(v[j + 3]) + b4.w;
      if (lrelu) {
        f0 = fmaxf(f0, slope * f0); f1 = fmaxf(f1, slope * f1);
        f2 = fmaxf(f2, slope * f2); f3 = fmaxf(f3, slope * f3);
      }
      pk[j / 2] = pack_bf16x2(f0, f1);
      pk[j / 2 + 1] = pack_bf16x2(f2, f3);
    }
    if (!full_n) {      // channel padding: stored as exact zeros
#pragma unroll
      for (int j = 0; j < CH; j += 2) {
        const int c = e.n0 + c0 + j;
        if (c >= e.co_real) pk[j / 2] = 0u;
        else if (c + 1 >= e.co_real) pk[j / 2] &= 0x0000ffffu;
      }
    }
    if (!valid) {
#pragma unroll
      for (int j = 0; j < CH / 2; ++j) pk[j] = 0u;
    }
    uint4* dst = reinterpret_cast<uint4*>(stg + row * PITCH + c0 * 2);
#pragma unroll
    for (int j = 0; j < CH / 8; ++j) dst[j] = make_uint4(pk[4 * j], pk[4 * j + 1], pk[4 * j + 2], pk[4 * j + 3]);
  }
  if (release_bar != nullptr) {   // the accumulator is in shared memory now: give the TMEM tile back
    tc_fence_before();
    __syncwarp();
    if (lane == 0) mbar_arrive(release_bar);
  }
  named_bar_sync(1, 128);
  // ---- phase 2: BatchNorm statistics, 8 channels x 16 rows per thread, then 8-way combine ------------
  if (e.stats != nullptr) {
    constexpr int G = BLOCK_N / 8;            // channel groups of 8
    constexpr int RG = 128 / G;               // row groups (threads per channel group)
    constexpr int RPT = 128 / RG;             // rows per thread
    const int g = et % G, rg = et / G;
    float s[8], ss[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) s[k] = ss[k] = 0.f;
#pragma unroll 4
    for (int r = rg * RPT; r < rg * RPT + RPT; ++r) {
      const uint4 u = *reinterpret_cast<const uint4*>(stg + r * PITCH + g * 16);
      const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float a = __uint_as_float(w[k] << 16), b = __uint_as_float(w[k] & 0xffff0000u);
        s[2 * k] += a; ss[2 * k] = fmaf(a, a, ss[2 * k]);
        s[2 * k + 1] += b; ss[2 * k + 1] = fmaf(b, b, ss[2 * k + 1]);
      }
    }
    // s_red[2][RG][BLOCK_N]
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      s_red[(0 * RG + rg) * BLOCK_N + g * 8 + k] = s[k];
      s_red[(1 * RG + rg) * BLOCK_N + g * 8 + k] = ss[k];
    }
    named_bar_sync(1, 128);
    for (int t = et; t < 2 * BLOCK_N; t += 128) {
      const int which = t / BLOCK_N, c = t % BLOCK_N;
      if (e.n0 + c < e.co_real) {
        float a = 0.f;
#pragma unroll
        for (int r = 0; r < RG; ++r) a += s_red[(which * RG + r) * BLOCK_N + c];
        atomicAdd(e.stats + which * e.Cn + e.n0 + c, (double)a);
      }
    }
  }
  // ---- phase 3: coalesced stores, LPR lanes cover one pixel's BLOCK_N channels (16 B each) ------------
  constexpr int LPR = BLOCK_N / 8;
  constexpr int RPI = 128 / LPR;
  const int sub = et % LPR, r_in = et / LPR;
#pragma unroll 2
  for (int r0 = 0; r0 < 128; r0 += RPI) {
    const int r = r0 + r_in;
    const int yy = e.oy0 + (r >> e.tw_log2), xx = e.ox0 + (r & tw_mask);
    if (yy < e.Hv && xx < e.Wv) {
      const uint4 val = *reinterpret_cast<const uint4*>(stg + r * PITCH + sub * 16);
      __nv_bfloat16* o = e.out + (int64_t)(yy * e.h_mul + e.h_off) * e.out_sH +
                         (int64_t)(xx * e.w_mul + e.w_off) * e.out_sW + e.n0 + sub * 8;
      *reinterpret_cast<uint4*>(o) = val;
    }
  }
  named_bar_sync(1, 128);   // staging (and s_red) may be reused
}

}  // namespace tc
}  // namespace p2p

// ---------------------------------------------------------------------------------------------------
// Split descriptors for tight MMA issue loops.  One thread issues every MMA, and a dependent chain of
// 64-bit descriptor arithmetic per instruction costs more cycles than a 128x128x16 MMA takes to execute
// (measured: ~250 cycles/MMA issue vs 64 cycles of tensor work).  The high word of a shared-memory
// descriptor (SBO, version, swizzle) is loop invariant; the low word is (address >> 4) | LBO << 16, so
// stepping through taps / K slices is one 32-bit add.
// ---------------------------------------------------------------------------------------------------
namespace p2p {
namespace tc {

__device__ __forceinline__ uint32_t desc_hi(uint32_t sbo_bytes, int row_bytes) {
  return ((sbo_bytes >> 4) & 0x3FFF) | (1u << 14) | (umma_layout_for_bytes(row_bytes) << 29);
}
__device__ __forceinline__ uint32_t desc_lo(uint32_t saddr, uint32_t lbo_bytes) {
  return ((saddr >> 4) & 0x3FFF) | (((lbo_bytes >> 4) & 0x3FFF) << 16);
}
__device__ __forceinline__ void umma_bf16_split(uint32_t tmem_d, uint32_t a_lo, uint32_t a_hi, uint32_t b_lo,
                                                uint32_t b_hi, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\t"
      "mov.b64 da, {%1, %2};\n\tmov.b64 db, {%3, %4};\n\t"
      "setp.ne.b32 p, %6, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %5, p;\n\t}\n" ::"r"(tmem_d),
      "r"(a_lo), "r"(a_hi), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(accumulate)
      : "memory");
}

}  // namespace tc
}  // namespace p2p
