// Halo-resident tcgen05 implicit-GEMM convolution for stride-1 problems (forward of the stride-1 layers and
// every data-gradient), sm_100a.  Reference layers: networks.py:404 (ConvLayer), :422 (UpsampleConvLayer),
// :776-784 (the stride-1 D convs), and autograd's conv backward-data for all of them.
//
// conv_tc.cu re-reads the input from L2 once per filter tap.  Here one TMA 4-D box brings the whole input
// halo of a 32x8-pixel output block (two 16x8 = 128-row UMMA tiles) into shared memory ONCE per 64-channel
// chunk, and every tap is just a different UMMA descriptor into that tile:
//     row(m, tap) = (m/8 + dh)*HW + (m%8 + dw)          start address = base + row*pitch, SBO = HW*pitch
// which works because tcgen05.mma applies the shared-memory swizzle to absolute address bits (probed on
// hardware: tools/probe/umma_shift_probe*.cu -- row-shifted starts and SBO != 8 rows give exact results for
// 32/64/128-byte rows).  Weights either stream through a ring, one [BLOCK_N x CIN_BLK] tile per (tap, chunk)
// feeding both accumulator tiles, or -- when the whole filter bank of the layer fits in ~100 KB (the 9x9
// layers, deconv2) -- are loaded once per CTA and stay resident, so the steady state moves only halos.
//
// Persistent kernel, one CTA per SM, 7 warps:  warp 0 = halo (A) TMA producer, warp 6 = weight (B) TMA
// producer, warp 1 = MMA issuer, warps 2-5 = epilogue.  TMEM holds two accumulator sets (2 tiles x BLOCK_N
// columns each), so the epilogue of block i overlaps the MMAs of block i+1.
#include "tc_common.cuh"

namespace p2p {

using namespace tc;

constexpr int kHaloMaxTaps = 81;

struct HaloTap {
  int32_t row_off16;  // ((dh - min_dh) * HW + (dw - min_dw)) * row_bytes / 16: descriptor-address step of the tap
  int32_t wk;         // first K index of this tap in the weight matrix
};

struct HaloParams {
  int n_taps, c_chunks, n_tiles;        // K loop; output-channel tiles
  int blocks_w, blocks_h, n_img;        // 32x8 pixel blocks per image
  int min_dh, min_dw, HW, HH;           // halo geometry (halo is HH x HW pixels)
  int rows;                             // output rows per block: 32, or 16..32 (even) in the swapped kernel
  int a_bytes;                          // bytes of one halo chunk buffer (1024-aligned)
  int nb;                               // weight ring depth (streaming) / number of resident tiles
  int resident;                         // 1: all n_taps * c_chunks weight tiles stay in shared memory
  EpiOut epi;                           // output addressing, bias handled separately
  int64_t out_sN;
  const float* bias;
  long long* clk;                       // debug clocks of the swapped kernel (nullptr in production)
  HaloTap taps[kHaloMaxTaps];
};

__device__ __forceinline__ unsigned long long gtimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

template <int BLOCK_N, int CIN_BLK>
__global__ void __launch_bounds__(224, 1)
    conv_halo_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                     const __grid_constant__ HaloParams p) {
  constexpr int RB = CIN_BLK * 2;                 // bytes per halo row (= swizzle span)
  constexpr int B_BYTES = BLOCK_N * RB;           // one weight tile
  constexpr int PITCH = BLOCK_N * 2 + 16;         // epilogue staging row pitch
  constexpr int NA = 2;
  constexpr uint32_t TMEM_COLS = 4 * BLOCK_N < 32 ? 32 : 4 * BLOCK_N;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sB = sA + NA * p.a_bytes;
  uint8_t* stg = sB + p.nb * B_BYTES;
  float* s_red = reinterpret_cast<float*>(stg + 128 * PITCH);
  __shared__ uint64_t a_full[NA], a_empty[NA], b_full[8], b_empty[8], t_full[2], t_empty[2];
  __shared__ uint32_t tmem_base_s;
  __shared__ __align__(16) float s_bias[BLOCK_N];

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int blocks_per_img = p.blocks_w * p.blocks_h;
  const int items_per_ntile = blocks_per_img * p.n_img;
  const int n_items = items_per_ntile * p.n_tiles;
  const int n_ring = p.resident ? 1 : p.nb;

  if (threadIdx.x == 0) {
    prefetch_tmap(&map_a);
    prefetch_tmap(&map_b);
    for (int s = 0; s < NA; ++s) { mbar_init(&a_full[s], 1); mbar_init(&a_empty[s], 1); }
    for (int s = 0; s < n_ring; ++s) { mbar_init(&b_full[s], 1); mbar_init(&b_empty[s], 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(&t_full[s], 1); mbar_init(&t_empty[s], 4); }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(&tmem_base_s, TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_s;

  // Role branches are WARP-uniform and the single issuing lane is chosen with elect.sync inside them: with a
  // per-thread branch (threadIdx.x == k) the compiler cannot keep TMA / MMA operands in uniform registers and
  // wraps every UTMALDG / UTCHMMA in a lane-serialising loop.
  if (warp == 0) {
    // ===== halo producer: one box per (item, channel chunk) =====
    if (elect_one()) {
      int ab = 0;
      uint32_t aph = 0;
      for (int it = blockIdx.x; it < n_items; it += gridDim.x) {
        const int r = it % items_per_ntile;
        const int img = r / blocks_per_img, br = r % blocks_per_img;
        const int oy0 = (br / p.blocks_w) * 32, ox0 = (br % p.blocks_w) * 8;
        for (int c = 0; c < p.c_chunks; ++c) {
          mbar_wait(&a_empty[ab], aph ^ 1);
          mbar_arrive_expect_tx(&a_full[ab], (uint32_t)(p.HH * p.HW * RB));
          tma_load_4d(sA + ab * p.a_bytes, &map_a, &a_full[ab], c * CIN_BLK, ox0 + p.min_dw, oy0 + p.min_dh, img);
          if (++ab == NA) { ab = 0; aph ^= 1; }
        }
      }
    }
    __syncwarp();
  } else if (warp == 6) {
    // ===== weight producer =====
    if (p.resident) {
      // the whole filter bank once: tile (c, t) lives at sB + (c * n_taps + t) * B_BYTES for the kernel's lifetime
      if (elect_one()) {
        mbar_arrive_expect_tx(&b_full[0], (uint32_t)(p.c_chunks * p.n_taps * B_BYTES));
        for (int c = 0; c < p.c_chunks; ++c)
          for (int t = 0; t < p.n_taps; ++t)
            tma_load_2d(sB + (c * p.n_taps + t) * B_BYTES, &map_b, &b_full[0], p.taps[t].wk + c * CIN_BLK, 0);
      }
      __syncwarp();
    } else {
      if (elect_one()) {
        int bs = 0;
        uint32_t bph = 0;
        for (int it = blockIdx.x; it < n_items; it += gridDim.x) {
          const int n0 = (it / items_per_ntile) * BLOCK_N;
          for (int c = 0; c < p.c_chunks; ++c)
            for (int t = 0; t < p.n_taps; ++t) {
              mbar_wait(&b_empty[bs], bph ^ 1);
              mbar_arrive_expect_tx(&b_full[bs], B_BYTES);
              tma_load_2d(sB + bs * B_BYTES, &map_b, &b_full[bs], p.taps[t].wk + c * CIN_BLK, n0);
              if (++bs == p.nb) { bs = 0; bph ^= 1; }
            }
        }
      }
      __syncwarp();
    }
  } else if (warp == 1) {
    // ===== MMA issuer (whole warp walks the loops, one elected lane issues) =====
    constexpr uint32_t idesc = umma_idesc_bf16(128, BLOCK_N, 0, 0);
    const uint32_t a_hi = desc_hi((uint32_t)(p.HW * RB), RB);     // 8-row groups of the tile are HW halo rows apart
    const uint32_t b_hi = desc_hi(8 * RB, RB);
    const uint32_t h_step = (uint32_t)(16 * p.HW * RB) >> 4;      // second tile: 16 halo rows further down
    int ab = 0, bs = 0, li = 0;
    uint32_t aph = 0, bph = 0;
    if (p.resident) mbar_wait(&b_full[0], 0);
    for (int it = blockIdx.x; it < n_items; it += gridDim.x, ++li) {
      const int set = li & 1;
      mbar_wait(&t_empty[set], ((li >> 1) & 1) ^ 1);
      tc_fence_after();
      const uint32_t acc = tmem_base + set * 2 * BLOCK_N;
      uint32_t first = 0;                                         // 0 for the first MMA of each accumulator
      for (int c = 0; c < p.c_chunks; ++c) {
        mbar_wait(&a_full[ab], aph);
        tc_fence_after();
        const uint32_t a_lo0 = desc_lo(smem_u32(sA + ab * p.a_bytes), 16);
        uint32_t b_lo = desc_lo(smem_u32(sB + (p.resident ? c * p.n_taps : 0) * B_BYTES), 16);
        if (p.resident) {
          // no barriers inside: the elected lane walks all taps of this chunk back to back
          if (elect_one()) {
#pragma unroll 3
            for (int t = 0; t < p.n_taps; ++t) {
              const uint32_t a_lo = a_lo0 + (uint32_t)p.taps[t].row_off16;
#pragma unroll
              for (int k = 0; k < CIN_BLK / 16; ++k) {
                umma_bf16_split(acc, a_lo + 2 * k, a_hi, b_lo + 2 * k, b_hi, idesc, first | (uint32_t)(t | k));
                umma_bf16_split(acc + BLOCK_N, a_lo + h_step + 2 * k, a_hi, b_lo + 2 * k, b_hi, idesc,
                                first | (uint32_t)(t | k));
              }
              b_lo += B_BYTES >> 4;
            }
          }
          __syncwarp();
          first = 1;
        } else {
          // one elected lane waits on the weight ring and issues; bs / bph are recomputed by every lane afterwards
          if (elect_one()) {
            int bs_l = bs;
            uint32_t bph_l = bph;
            for (int t = 0; t < p.n_taps; ++t) {
              mbar_wait(&b_full[bs_l], bph_l);
              tc_fence_after();
              const uint32_t b_lo_t = desc_lo(smem_u32(sB + bs_l * B_BYTES), 16);
              const uint32_t a_lo = a_lo0 + (uint32_t)p.taps[t].row_off16;
#pragma unroll
              for (int k = 0; k < CIN_BLK / 16; ++k) {
                umma_bf16_split(acc, a_lo + 2 * k, a_hi, b_lo_t + 2 * k, b_hi, idesc, first | (uint32_t)(t | k));
                umma_bf16_split(acc + BLOCK_N, a_lo + h_step + 2 * k, a_hi, b_lo_t + 2 * k, b_hi, idesc,
                                first | (uint32_t)(t | k));
              }
              umma_commit(&b_empty[bs_l]);
              if (++bs_l == p.nb) { bs_l = 0; bph_l ^= 1; }
            }
          }
          __syncwarp();
          first = 1;
          for (int t = 0; t < p.n_taps; ++t)
            if (++bs == p.nb) { bs = 0; bph ^= 1; }
        }
        if (elect_one()) umma_commit(&a_empty[ab]);
        __syncwarp();
        if (++ab == NA) { ab = 0; aph ^= 1; }
      }
      if (elect_one()) umma_commit(&t_full[set]);
      __syncwarp();
    }
  } else if (warp >= 2 && warp <= 5) {
    // ===== epilogue =====
    const int q = warp & 3;
    const int et = threadIdx.x - 64;        // 0..127
    int li = 0;
    int cur_n0 = -1;
    EpiOut e = p.epi;
    e.tw_log2 = 3;
    for (int it = blockIdx.x; it < n_items; it += gridDim.x, ++li) {
      const int set = li & 1;
      const int n0 = (it / items_per_ntile) * BLOCK_N;
      const int r = it % items_per_ntile;
      const int img = r / blocks_per_img, br = r % blocks_per_img;
      const int oy0 = (br / p.blocks_w) * 32, ox0 = (br % p.blocks_w) * 8;
      if (n0 != cur_n0) {
        named_bar_sync(1, 128);
        for (int i = et; i < BLOCK_N; i += 128)
          s_bias[i] = (p.bias != nullptr && n0 + i < e.co_real) ? p.bias[n0 + i] : 0.f;
        named_bar_sync(1, 128);
        cur_n0 = n0;
      }
      e.n0 = n0;
      e.out = p.epi.out + (int64_t)img * p.out_sN;
      e.ox0 = ox0;
      mbar_wait(&t_full[set], (li >> 1) & 1);
      tc_fence_after();
#pragma unroll 1
      for (int h = 0; h < 2; ++h) {
        e.oy0 = oy0 + h * 16;
        epilogue_tile<BLOCK_N>(tmem_base + (uint32_t)(set * 2 * BLOCK_N + h * BLOCK_N), q, lane, et, stg, s_bias, s_red, e,
                               h == 1 ? &t_empty[set] : nullptr);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, TMEM_COLS);
}

// ---------------------------------------------------------------------------------------------------
// Operand-swapped variant for 128-wide output-channel tiles with streamed weights (the residual 128->128
// convolutions and every data-gradient into 128 channels).
//
// The kernel above is bound by shared-memory bandwidth, not by TMA or the tensor pipe
// (profiles/r01d_halo_stall_clocks.md): per 256-pixel item its 144 M=128 x N=128 MMAs read 144 x 8 KB of
// operands, TMA writes 382 KB and the epilogue moves 192 KB through the staging buffer -- 1.73 MB at
// 128 B/cycle is 13.5 k cycles against a tensor floor of 9.2 k.  Here the weight tile is the A operand
// (M = 128 output channels) and BOTH 16x8-pixel tiles of the block are one N = 256 B operand, read through the
// same shifted halo descriptor (32 row groups HW rows apart): 72 MMAs x 12 KB = 0.86 MB.  The accumulator is
// then transposed -- TMEM lane = output channel, column = pixel -- so an epilogue thread owns one channel:
// bias and the BatchNorm sums are thread-local (no shared memory at all), and neighbouring lanes exchange one
// value per pixel pair so that every lane stores 4 bytes and a warp covers 64 contiguous bytes of two pixels.
// ---------------------------------------------------------------------------------------------------
// PROBE = true compiles in the clock64 / %globaltimer stamps read by tools/halo_clocks.py (a separate instantiation:
// the production kernel contains none of it).
template <int CIN_BLK, bool PROBE>
__global__ void __launch_bounds__(352, 1)
    conv_halo_swap_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                          const __grid_constant__ HaloParams p) {
  constexpr int RB = CIN_BLK * 2;
  constexpr int W_BYTES = 128 * RB;               // one [128 output channels x CIN_BLK] weight tile
  constexpr int NA = 2;
  constexpr uint32_t TMEM_COLS = 512;             // two accumulator sets of 256 pixel columns
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;                             // halo chunks
  uint8_t* sB = sA + NA * p.a_bytes;              // weight ring
  __shared__ uint64_t a_full[NA], a_empty[NA], b_full[8], b_empty[8], t_full[2], t_empty[2];
  __shared__ uint32_t tmem_base_s;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const bool probe = PROBE && p.clk != nullptr;
#define P2P_CLK() (PROBE ? clock64() : 0ll)
  if (probe && threadIdx.x == 0) p.clk[blockIdx.x * 16 + 8] = (long long)gtimer();
  const int blocks_per_img = p.blocks_w * p.blocks_h;
  const int items_per_ntile = blocks_per_img * p.n_img;
  const int n_items = items_per_ntile * p.n_tiles;

  if (threadIdx.x == 0) {
    prefetch_tmap(&map_a);
    prefetch_tmap(&map_b);
    for (int s = 0; s < NA; ++s) { mbar_init(&a_full[s], 1); mbar_init(&a_empty[s], 1); }
    for (int s = 0; s < p.nb; ++s) { mbar_init(&b_full[s], 1); mbar_init(&b_empty[s], 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(&t_full[s], 1); mbar_init(&t_empty[s], 8); }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(&tmem_base_s, TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_s;

  if (warp == 0) {
    // ===== halo producer =====
    if (elect_one()) {
      int ab = 0;
      uint32_t aph = 0;
      for (int it = blockIdx.x; it < n_items; it += gridDim.x) {
        const int r = it % items_per_ntile;
        const int img = r / blocks_per_img, br = r % blocks_per_img;
        const int oy0 = (br / p.blocks_w) * p.rows, ox0 = (br % p.blocks_w) * 8;
        for (int c = 0; c < p.c_chunks; ++c) {
          mbar_wait(&a_empty[ab], aph ^ 1);
          mbar_arrive_expect_tx(&a_full[ab], (uint32_t)(p.HH * p.HW * RB));
          tma_load_4d(sA + ab * p.a_bytes, &map_a, &a_full[ab], c * CIN_BLK, ox0 + p.min_dw, oy0 + p.min_dh, img);
          if (++ab == NA) { ab = 0; aph ^= 1; }
        }
      }
    }
    __syncwarp();
  } else if (warp == 10) {
    // ===== weight producer =====
    if (elect_one()) {
      int bs = 0;
      uint32_t bph = 0;
      for (int it = blockIdx.x; it < n_items; it += gridDim.x) {
        const int n0 = (it / items_per_ntile) * 128;
        for (int c = 0; c < p.c_chunks; ++c)
          for (int t = 0; t < p.n_taps; ++t) {
            mbar_wait(&b_empty[bs], bph ^ 1);
            mbar_arrive_expect_tx(&b_full[bs], W_BYTES);
            tma_load_2d(sB + bs * W_BYTES, &map_b, &b_full[bs], p.taps[t].wk + c * CIN_BLK, n0);
            if (++bs == p.nb) { bs = 0; bph ^= 1; }
          }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ===== MMA issuer: D[channel, pixel] += W[channel, k] * X[pixel, k] =====
    const uint32_t idesc = umma_idesc_bf16(128, 8 * p.rows, 0, 0);   // N = pixels of one block (multiple of 16)
    const uint32_t x_hi = desc_hi((uint32_t)(p.HW * RB), RB);     // pixel row groups are HW halo rows apart
    const uint32_t w_hi = desc_hi(8 * RB, RB);
    int ab = 0, bs = 0, li = 0;
    uint32_t aph = 0, bph = 0;
    const long long c_tot = P2P_CLK();
    long long c_te = 0, c_af = 0, c_t0;
    if (probe && lane == 0) p.clk[blockIdx.x * 16 + 9] = (long long)gtimer();
    for (int it = blockIdx.x; it < n_items; it += gridDim.x, ++li) {
      const int set = li & 1;
      c_t0 = P2P_CLK();
      mbar_wait(&t_empty[set], ((li >> 1) & 1) ^ 1);
      c_te += P2P_CLK() - c_t0;
      tc_fence_after();
      const uint32_t acc = tmem_base + set * 256;
      uint32_t first = 0;
      for (int c = 0; c < p.c_chunks; ++c) {
        c_t0 = P2P_CLK();
        mbar_wait(&a_full[ab], aph);
        c_af += P2P_CLK() - c_t0;
        tc_fence_after();
        const uint32_t x_lo0 = desc_lo(smem_u32(sA + ab * p.a_bytes), 16);
        if (elect_one()) {
          int bs_l = bs;
          uint32_t bph_l = bph;
          long long c_bf = 0;
          for (int t = 0; t < p.n_taps; ++t) {
            const long long c_b0 = P2P_CLK();
            mbar_wait(&b_full[bs_l], bph_l);
            c_bf += P2P_CLK() - c_b0;
            tc_fence_after();
            const uint32_t w_lo = desc_lo(smem_u32(sB + bs_l * W_BYTES), 16);
            const uint32_t x_lo = x_lo0 + (uint32_t)p.taps[t].row_off16;
#pragma unroll
            for (int k = 0; k < CIN_BLK / 16; ++k)
              umma_bf16_split(acc, w_lo + 2 * k, w_hi, x_lo + 2 * k, x_hi, idesc, first | (uint32_t)(t | k));
            umma_commit(&b_empty[bs_l]);
            if (++bs_l == p.nb) { bs_l = 0; bph_l ^= 1; }
          }
          if (probe)
            atomicAdd(reinterpret_cast<unsigned long long*>(p.clk + blockIdx.x * 16 + 3), (unsigned long long)c_bf);
        }
        __syncwarp();
        first = 1;
        for (int t = 0; t < p.n_taps; ++t)
          if (++bs == p.nb) { bs = 0; bph ^= 1; }
        if (elect_one()) umma_commit(&a_empty[ab]);
        __syncwarp();
        if (++ab == NA) { ab = 0; aph ^= 1; }
      }
      if (elect_one()) umma_commit(&t_full[set]);
      __syncwarp();
    }
    if (probe && lane == 0) {
      p.clk[blockIdx.x * 16 + 0] = P2P_CLK() - c_tot;
      p.clk[blockIdx.x * 16 + 1] = c_te;
      p.clk[blockIdx.x * 16 + 2] = c_af;
      p.clk[blockIdx.x * 16 + 10] = (long long)gtimer();
    }
  } else if (warp >= 2 && warp <= 9) {
    // ===== epilogue (8 warps): this thread owns output channel n0 + ch for one half (16 rows) of the block =====
    const int q = warp & 3;                     // TMEM lane quarter this warp may read
    const int half = (warp - 2) >> 2;           // columns [128 * half, 128 * half + 128) = block rows 16*half ..
    const int ch = q * 32 + lane;
    const bool lrelu = p.epi.act == P2P_ACT_LRELU || p.epi.act == P2P_ACT_RELU;   // max(f, slope * f)
    const float slope = p.epi.act == P2P_ACT_LRELU ? 0.2f : 0.f;
    const bool even = (lane & 1) == 0;
    // lane pairs trade halves so that every lane owns one pixel x two channels: odd lanes first swap their own
    // halves (sel_own), then each lane keeps its low half and takes the neighbour's high half (sel_word)
    const uint32_t sel_own = even ? 0x3210u : 0x1032u;
    const uint32_t sel_word = even ? 0x7610u : 0x1076u;
    const int64_t sx = (int64_t)p.epi.w_mul * p.epi.out_sW;      // elements between horizontally adjacent pixels
    int li = 0, cur_n0 = -1;
    float bias = 0.f;
    bool real = false;
    double d_s = 0.0, d_ss = 0.0;
    for (int it = blockIdx.x; it < n_items; it += gridDim.x, ++li) {
      const int set = li & 1;
      const int n0 = (it / items_per_ntile) * 128;
      const int r = it % items_per_ntile;
      const int img = r / blocks_per_img, br = r % blocks_per_img;
      const int oyb = (br / p.blocks_w) * p.rows;
      const int oy0 = oyb + half * 16, ox0 = (br % p.blocks_w) * 8;
      const int h_lim = min(p.epi.Hv, oyb + p.rows);            // rows of this block that exist
      const int n_cols = 8 * min(16, p.rows - half * 16);       // accumulator columns of this half (>= 0)
      if (n0 != cur_n0) {
        if (cur_n0 >= 0 && real && p.epi.stats != nullptr) {
          atomicAdd(p.epi.stats + cur_n0 + ch, d_s);
          atomicAdd(p.epi.stats + p.epi.Cn + cur_n0 + ch, d_ss);
        }
        d_s = d_ss = 0.0;
        real = n0 + ch < p.epi.co_real;
        bias = (p.bias != nullptr && real) ? p.bias[n0 + ch] : 0.f;
        cur_n0 = n0;
      }
      // the even lane stores channels (ch, ch+1) of the even pixel of a pair, the odd lane (ch-1, ch) of the odd one
      __nv_bfloat16* obase = p.epi.out + (int64_t)img * p.out_sN + n0 + (even ? ch : ch - 1) +
                             (int64_t)(ox0 * p.epi.w_mul + p.epi.w_off) * p.epi.out_sW + (even ? 0 : sx);
      const bool all_real_x = __all_sync(0xffffffffu, real && ox0 + 8 <= p.epi.Wv);
      const long long c_e0 = P2P_CLK();
      mbar_wait(&t_full[set], (li >> 1) & 1);
      const long long c_e1 = P2P_CLK();
      tc_fence_after();
      float s = 0.f, ss = 0.f;
      if (n_cols <= 0) {                         // blocks of 16 rows: the upper half has nothing to read
        __syncwarp();
        if (lane == 0) mbar_arrive(&t_empty[set]);
      }
#pragma unroll 1
      for (int j0 = 0; j0 < n_cols; j0 += 32) {
        uint32_t v[32];
        tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(set * 256 + half * 128 + j0), v);
        tmem_ld_wait();
        if (j0 + 32 >= n_cols) {                // last read of this accumulator set: hand it back to the MMA warp
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&t_empty[set]);
        }
        const int yy0 = oy0 + (j0 >> 3);        // the 32 columns are 4 block rows of 8 pixels
        if (all_real_x && yy0 + 4 <= h_lim) {
#pragma unroll
          for (int rr = 0; rr < 4; ++rr) {
            __nv_bfloat16* orow = obase + (int64_t)((yy0 + rr) * p.epi.h_mul + p.epi.h_off) * p.epi.out_sH;
#pragma unroll
            for (int x = 0; x < 8; x += 2) {
              const int j = rr * 8 + x;
              float f0 = __uint_as_float(v[j]) + bias, f1 = __uint_as_float(v[j + 1]) + bias;
              if (lrelu) { f0 = fmaxf(f0, slope * f0); f1 = fmaxf(f1, slope * f1); }
              const uint32_t pk = pack_bf16x2(f0, f1);    // (pixel j, pixel j+1) of my channel, rounded as stored
              const float r0 = __uint_as_float(pk << 16), r1 = __uint_as_float(pk & 0xffff0000u);
              s += r0 + r1;
              ss = fmaf(r0, r0, fmaf(r1, r1, ss));
              const uint32_t own = __byte_perm(pk, 0u, sel_own);
              const uint32_t got = __shfl_xor_sync(0xffffffffu, own, 1);
              *reinterpret_cast<uint32_t*>(orow + x * sx) = __byte_perm(own, got, sel_word);
            }
          }
        } else {
#pragma unroll
          for (int j = 0; j < 32; j += 2) {
            const int yy = yy0 + (j >> 3), x = j & 7;
            const bool row_ok = yy < h_lim;
            const bool ok0 = row_ok && ox0 + x < p.epi.Wv, ok1 = row_ok && ox0 + x + 1 < p.epi.Wv;
            float f0 = __uint_as_float(v[j]) + bias, f1 = __uint_as_float(v[j + 1]) + bias;
            if (lrelu) { f0 = fmaxf(f0, slope * f0); f1 = fmaxf(f1, slope * f1); }
            if (!real || !ok0) f0 = 0.f;        // channel padding and positions outside the map are exact zeros
            if (!real || !ok1) f1 = 0.f;
            const uint32_t pk = pack_bf16x2(f0, f1);
            const float r0 = __uint_as_float(pk << 16), r1 = __uint_as_float(pk & 0xffff0000u);
            s += r0 + r1;
            ss = fmaf(r0, r0, fmaf(r1, r1, ss));
            const uint32_t own = __byte_perm(pk, 0u, sel_own);
            const uint32_t got = __shfl_xor_sync(0xffffffffu, own, 1);
            if (even ? ok0 : ok1)
              *reinterpret_cast<uint32_t*>(obase + (int64_t)(yy * p.epi.h_mul + p.epi.h_off) * p.epi.out_sH + x * sx) =
                  __byte_perm(own, got, sel_word);
          }
        }
      }
      d_s += (double)s;
      d_ss += (double)ss;
      if (probe && threadIdx.x == 64) {
        p.clk[blockIdx.x * 16 + 5] += c_e1 - c_e0;
        p.clk[blockIdx.x * 16 + 6] += P2P_CLK() - c_e1;
        p.clk[blockIdx.x * 16 + 11] = (long long)gtimer();
      }
    }
    if (cur_n0 >= 0 && real && p.epi.stats != nullptr) {
      atomicAdd(p.epi.stats + cur_n0 + ch, d_s);
      atomicAdd(p.epi.stats + p.epi.Cn + cur_n0 + ch, d_ss);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, TMEM_COLS);
  if (probe && threadIdx.x == 32) p.clk[blockIdx.x * 16 + 12] = (long long)gtimer();
#undef P2P_CLK
}

// ---------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------
// Debug hook (deliberately not in include/p2p_b200.h): tools/halo_clocks.py hands in a [148 x 16] int64 device
// buffer and the swapped kernel records, per CTA, clock64 sums of its mbarrier waits and %globaltimer stamps.
static long long* g_halo_clk = nullptr;
extern "C" void p2p_debug_halo_clocks(void* buf) { g_halo_clk = static_cast<long long*>(buf); }

template <int CIN_BLK>
static int launch_halo_swap(const CUtensorMap& ma, const CUtensorMap& mb, const HaloParams& p, int grid, int smem,
                            cudaStream_t st) {
  trace_kernel("conv_halo_swap<%d> rows=%d taps=%d chunks=%d nb=%d blocks=%dx%d", CIN_BLK, p.rows, p.n_taps, p.c_chunks,
               p.nb, p.blocks_h, p.blocks_w);
  if (p.clk != nullptr) {           // tools/halo_clocks.py only
    static SmemOptIn optin_probe;
    P2P_CUDA(ensure_dyn_smem(conv_halo_swap_kernel<CIN_BLK, true>, optin_probe, smem));
    conv_halo_swap_kernel<CIN_BLK, true><<<grid, 352, smem, st>>>(ma, mb, p);
    P2P_LAUNCH_CHECK();
  }
  static SmemOptIn optin;
  P2P_CUDA(ensure_dyn_smem(conv_halo_swap_kernel<CIN_BLK, false>, optin, smem));
  conv_halo_swap_kernel<CIN_BLK, false><<<grid, 352, smem, st>>>(ma, mb, p);
  P2P_LAUNCH_CHECK();
}

template <int BLOCK_N, int CIN_BLK>
static int launch_halo(const CUtensorMap& ma, const CUtensorMap& mb, const HaloParams& p, int grid, int smem,
                       cudaStream_t st) {
  static SmemOptIn optin;
  P2P_CUDA(ensure_dyn_smem(conv_halo_kernel<BLOCK_N, CIN_BLK>, optin, smem));
  trace_kernel("conv_halo<%d,%d> resident=%d taps=%d chunks=%d nb=%d", BLOCK_N, CIN_BLK, p.resident, p.n_taps, p.c_chunks,
               p.nb);
  conv_halo_kernel<BLOCK_N, CIN_BLK><<<grid, 224, smem, st>>>(ma, mb, p);
  P2P_LAUNCH_CHECK();
}

template <int CIN_BLK>
static int launch_halo_n(int block_n, const CUtensorMap& ma, const CUtensorMap& mb, const HaloParams& p, int grid,
                         int smem, cudaStream_t st) {
  switch (block_n) {
    case 16: return launch_halo<16, CIN_BLK>(ma, mb, p, grid, smem, st);
    case 32: return launch_halo<32, CIN_BLK>(ma, mb, p, grid, smem, st);
    case 64: return launch_halo<64, CIN_BLK>(ma, mb, p, grid, smem, st);
    default: return launch_halo<128, CIN_BLK>(ma, mb, p, grid, smem, st);
  }
}

// A stride-1 implicit-GEMM problem.  taps: (dh, dw, wk) triples.  Returns -2 when the shape does not fit or
// the per-tap kernel of conv_tc.cu is the better choice (the caller then uses that one).
int run_halo_problem(const void* a_src, int N, int Hs, int Ws, int Ck, const void* b_mat, int Cn, int64_t ktot,
                     const int* dh, const int* dw, const int* wk, int n_taps, int Hv, int Wv, __nv_bfloat16* out,
                     int64_t out_sN, int64_t out_sH, int64_t out_sW, int h_mul, int h_off, int w_mul, int w_off,
                     int co_real, int act, const float* bias, double* stats, bool force, cudaStream_t st) {
  if (n_taps > kHaloMaxTaps) return -2;
  const int cin_blk = Ck >= 64 ? 64 : Ck;
  const int block_n = Cn >= 128 ? 128 : Cn;
  const int rb = cin_blk * 2;
  int mn_h = dh[0], mx_h = dh[0], mn_w = dw[0], mx_w = dw[0];
  for (int t = 1; t < n_taps; ++t) {
    mn_h = dh[t] < mn_h ? dh[t] : mn_h; mx_h = dh[t] > mx_h ? dh[t] : mx_h;
    mn_w = dw[t] < mn_w ? dw[t] : mn_w; mx_w = dw[t] > mx_w ? dw[t] : mx_w;
  }
  HaloParams p{};
  p.HW = 8 + (mx_w - mn_w);
  p.HH = 32 + (mx_h - mn_h);
  if (p.HW > 256 || p.HH > 256) return -2;
  p.min_dh = mn_h;
  p.min_dw = mn_w;
  p.a_bytes = ((p.HH * p.HW * rb + 1023) / 1024) * 1024;
  p.c_chunks = Ck / cin_blk;
  p.n_tiles = Cn / block_n;
  const int b_bytes = block_n * rb;
  const int stg_bytes = 128 * (block_n * 2 + 16) + 8192;   // staging + statistics scratch
  const int budget = 225 * 1024 - 2 * p.a_bytes - stg_bytes - 1024;
  const int bank_tiles = n_taps * p.c_chunks;
  p.resident = (p.n_tiles == 1 && bank_tiles * b_bytes <= budget && bank_tiles * b_bytes <= 112 * 1024) ? 1 : 0;
  // 128-wide tiles with streamed weights run on the operand-swapped kernel (no staging buffer there)
  static const bool no_swap = getenv("P2P_NO_SWAP") != nullptr && atoi(getenv("P2P_NO_SWAP")) != 0;
  const bool swap = !p.resident && block_n == 128 && !no_swap;
  p.rows = 32;
  if (swap) {
    // N = 8 * rows pixels per MMA: split the map height evenly over its row blocks (66 rows -> 3 x 22, not 3 x 32)
    const int nbh = (Hv + 31) / 32;
    const int rows = 2 * ((Hv + 2 * nbh - 1) / (2 * nbh));
    p.rows = rows < 16 ? 16 : rows;
  }
  if (!force) {
    // Measured on B200 (tools/bench_convs.py, B=16).  Resident weights always beat the per-tap kernel.  With
    // streamed weights the plain kernel wins only for narrow tiles (Cn <= 64); the operand-swapped kernel wins
    // for every 128-multiple Cn (residual fwd 29.6 vs 35.6 us, deconv3 dgrad 52 vs 78, D 256->512 fwd 98 vs 118,
    // 18x18 maps 46 vs 62) except the parity phases of a stride-2 data-gradient, which are four small launches
    // (D 128<-256: 73 vs 56 us).
    if (!p.resident && Cn > 64 && !(swap && h_mul == 1 && w_mul == 1)) return -2;
    const int64_t blocks = (int64_t)((Wv + 7) / 8) * ((Hv + p.rows - 1) / p.rows) * N;
    if (blocks * 8 * p.rows > 2 * (int64_t)N * Hv * Wv) return -2;   // small maps: too many padded pixels per block
  }
  int ring_budget = budget;
  if (swap) {
    p.HH = p.rows + (mx_h - mn_h);
    p.a_bytes = ((p.HH * p.HW * rb + 1023) / 1024) * 1024;
    ring_budget = 225 * 1024 - 2 * p.a_bytes - 1024;
  }
  int nb;
  if (p.resident) {
    nb = bank_tiles;
  } else {
    nb = ring_budget / b_bytes;
    nb = nb > 8 ? 8 : nb;
    if (nb < 2) return -2;
  }
  p.nb = nb;
  const int smem = 2 * p.a_bytes + nb * b_bytes + (swap ? 0 : stg_bytes) + 1024;
  if (smem > 226 * 1024) return -2;
  p.n_taps = n_taps;
  for (int t = 0; t < n_taps; ++t) {
    const int ro = (dh[t] - mn_h) * p.HW + (dw[t] - mn_w);
    p.taps[t].row_off16 = ro * rb / 16;
    p.taps[t].wk = wk[t];
  }
  p.blocks_w = (Wv + 7) / 8;
  p.blocks_h = (Hv + p.rows - 1) / p.rows;
  p.n_img = N;
  p.out_sN = out_sN;
  p.bias = bias;
  p.epi.out = out;
  p.epi.out_sH = out_sH; p.epi.out_sW = out_sW;
  p.epi.h_mul = h_mul; p.epi.h_off = h_off; p.epi.w_mul = w_mul; p.epi.w_off = w_off;
  p.epi.Hv = Hv; p.epi.Wv = Wv;
  p.epi.co_real = co_real; p.epi.Cn = Cn; p.epi.act = act; p.epi.stats = stats;
  p.clk = g_halo_clk;

  CUtensorMap ma, mb;
  {
    const uint64_t dims[4] = {(uint64_t)Ck, (uint64_t)Ws, (uint64_t)Hs, (uint64_t)N};
    const uint64_t strides[3] = {(uint64_t)Ck * 2, (uint64_t)Ws * Ck * 2, (uint64_t)Hs * Ws * Ck * 2};
    const uint32_t box[4] = {(uint32_t)cin_blk, (uint32_t)p.HW, (uint32_t)p.HH, 1};
    const uint32_t es[4] = {1, 1, 1, 1};
    if (int e = make_map_bf16(&ma, a_src, 4, dims, strides, box, es, rb)) return e;
  }
  {
    const uint64_t dims[2] = {(uint64_t)ktot, (uint64_t)Cn};
    const uint64_t strides[1] = {(uint64_t)ktot * 2};
    const uint32_t box[2] = {(uint32_t)cin_blk, (uint32_t)block_n};
    const uint32_t es[2] = {1, 1};
    if (int e = make_map_bf16(&mb, b_mat, 2, dims, strides, box, es, rb)) return e;
  }
  const int n_items = p.blocks_w * p.blocks_h * N * p.n_tiles;
  const int grid = n_items < kSMs ? n_items : kSMs;
  if (swap) {
    switch (cin_blk) {
      case 16: return launch_halo_swap<16>(ma, mb, p, grid, smem, st);
      case 32: return launch_halo_swap<32>(ma, mb, p, grid, smem, st);
      default: return launch_halo_swap<64>(ma, mb, p, grid, smem, st);
    }
  }
  switch (cin_blk) {
    case 16: return launch_halo_n<16>(block_n, ma, mb, p, grid, smem, st);
    case 32: return launch_halo_n<32>(block_n, ma, mb, p, grid, smem, st);
    default: return launch_halo_n<64>(block_n, ma, mb, p, grid, smem, st);
  }
}

}  // namespace p2p
