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
  int a_bytes;                          // bytes of one halo chunk buffer (1024-aligned)
  int nb;                               // weight ring depth (streaming) / number of resident tiles
  int resident;                         // 1: all n_taps * c_chunks weight tiles stay in shared memory
  EpiOut epi;                           // output addressing, bias handled separately
  int64_t out_sN;
  const float* bias;
  HaloTap taps[kHaloMaxTaps];
};

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
// host side
// ---------------------------------------------------------------------------------------------------
template <int BLOCK_N, int CIN_BLK>
static int launch_halo(const CUtensorMap& ma, const CUtensorMap& mb, const HaloParams& p, int grid, int smem,
                       cudaStream_t st) {
  static int attr_smem = 0;
  if (smem > attr_smem) {
    P2P_CUDA(cudaFuncSetAttribute(conv_halo_kernel<BLOCK_N, CIN_BLK>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    attr_smem = smem;
  }
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
  if (!force) {
    // Measured on B200 (tools/bench_convs.py): with streamed weights this kernel beats the per-tap kernel only
    // where the output tile is narrow (Cn <= 64) or a single 128-wide tile reads at most two channel chunks
    // (residual 128->128: 33.0 vs 35.6 us, its dgrad 38.8 vs 41.8; deconv3 dgrad 62.8 vs 77.9); D's 256->512
    // layers lose (140 vs 118 us).  With resident weights it always wins.
    const int64_t blocks = (int64_t)((Wv + 7) / 8) * ((Hv + 31) / 32) * N;
    if (!p.resident && Cn > 64 && !(p.n_tiles == 1 && p.c_chunks <= 2)) return -2;
    if (blocks * 256 > 2 * (int64_t)N * Hv * Wv) return -2;    // small maps: too many padded pixels per 32x8 block
  }
  int nb;
  if (p.resident) {
    nb = bank_tiles;
  } else {
    nb = budget / b_bytes;
    nb = nb > 8 ? 8 : nb;
    if (nb < 2) return -2;
  }
  p.nb = nb;
  const int smem = 2 * p.a_bytes + nb * b_bytes + stg_bytes + 1024;
  if (smem > 226 * 1024) return -2;
  p.n_taps = n_taps;
  for (int t = 0; t < n_taps; ++t) {
    const int ro = (dh[t] - mn_h) * p.HW + (dw[t] - mn_w);
    p.taps[t].row_off16 = ro * rb / 16;
    p.taps[t].wk = wk[t];
  }
  p.blocks_w = (Wv + 7) / 8;
  p.blocks_h = (Hv + 31) / 32;
  p.n_img = N;
  p.out_sN = out_sN;
  p.bias = bias;
  p.epi.out = out;
  p.epi.out_sH = out_sH; p.epi.out_sW = out_sW;
  p.epi.h_mul = h_mul; p.epi.h_off = h_off; p.epi.w_mul = w_mul; p.epi.w_off = w_off;
  p.epi.Hv = Hv; p.epi.Wv = Wv;
  p.epi.co_real = co_real; p.epi.Cn = Cn; p.epi.act = act; p.epi.stats = stats;

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
  switch (cin_blk) {
    case 16: return launch_halo_n<16>(block_n, ma, mb, p, grid, smem, st);
    case 32: return launch_halo_n<32>(block_n, ma, mb, p, grid, smem, st);
    default: return launch_halo_n<64>(block_n, ma, mb, p, grid, smem, st);
  }
}

}  // namespace p2p
