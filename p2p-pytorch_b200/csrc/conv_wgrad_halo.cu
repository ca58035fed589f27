// Halo-resident tcgen05 weight-gradient kernel for stride-1 layers with Ci <= 128 (sm_100a):
//     dW[tap][ci][co] += sum over output pixels  x[p + tap][ci] * dy[p][co]
// (backward of nn.Conv2d w.r.t. its weight for the generator's stride-1 layers, reference networks.py:404,
// :422 through autograd: conv1, the 18 residual convs, deconv3/2/1.)
//
// conv_wgrad_tc.cu loads one TMA box of x per filter tap.  Here the x halo of an 8x16-pixel block is loaded
// ONCE per stage and every tap reads it through its own UMMA descriptor (tcgen05.mma applies the swizzle to
// absolute shared-memory address bits -- tools/probe/umma_shift_probe2.cu, mode 3):
//   * operands are MN-major: shared-memory rows are pixels (the reduction index), a row holds the channels;
//     one MMA reduces over the 16 pixels of one block row (K = 16);
//   * one M tile (128 accumulator lanes) is a chain of "atoms" of CIN_BLK channels.  For Ci = 128 the two
//     atoms are the two 64-channel halves of one tap (LBO = distance between the two halo buffers).  For
//     Ci <= 64 the atoms are CONSECUTIVE TAPS reading overlapping windows of the same halo: along kx the
//     next atom starts one halo row later (LBO = one row), along ky one halo line later (LBO = HW rows);
//   * a CTA owns up to 512 TMEM columns of accumulators (MT tiles x N = Co columns), walks its share of the
//     pixel blocks (split-K), then the 4 epilogue warps red.global.add.v4.f32 the tiles into dW.
// Per MMA the tensor core spends ~69 cycles for any N <= 128 (tools/probe/umma_rate_probe.cu), so the layer
// floor is (#M tiles x #pixels / 16 / #SMs) x 69 cycles; the point of the halo is that TMA traffic no longer
// caps the kernel below that floor (the per-tap kernel moves ~77 B/cycle/SM, TMA sustains ~32).
#include "tc_common.cuh"

namespace p2p {

using namespace tc;

constexpr int kWhMaxTiles = 24;
constexpr int kWhRows = 8, kWhCols = 16;   // pixel block of one stage

struct WhTile {
  uint32_t lo_const;   // descriptor low-word constant: (LBO >> 4) << 16 | (row offset of the first atom * row_bytes) >> 4
  int16_t tap[8];      // flattened tap index of each atom, -1 = padding atom (result discarded)
  int16_t c0[8];       // first input channel of each atom
};

struct WhParams {
  int n_tiles_m, mt;                 // M tiles in total / per CTA
  int cin_blk, n_chunks;             // channels per atom, halo buffers per stage (Ci / 64, at least 1)
  int HWx, HHx, x_bytes;             // halo geometry and bytes of one halo buffer (1024-aligned)
  int min_dh, min_dw;
  int co_blk, n_cols, b_boxes;       // dy: channels per box, UMMA N, boxes per stage
  int stages, stage_bytes;
  int blocks_w, blocks_h, n_img, blocks_per_split;
  int Ci, Co;
  // "ky on the B side" mode (Ci, Co <= 64): the M tiles chain only the kx taps; the N dimension of one MMA is
  // (g_ky filter rows) x Co, read from a dy tile with a (g_ky - 1)-line halo above the block -- atom j of the B
  // operand is the same 16-pixel line of dy one line further down (LBO = one line), i.e. filter row ky_hi - j.
  // blockIdx.y then selects the group of filter rows instead of an output-channel tile.
  int ky_b, g_ky, kh, kw, pad, HHy;
  float* dw;
  WhTile tiles[kWhMaxTiles];
};

__global__ void __launch_bounds__(192, 1)
    conv_wgrad_halo_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_dy,
                           const __grid_constant__ WhParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t full_bar[8], empty_bar[8], tmem_full_bar;
  __shared__ uint32_t tmem_base_s;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int rbx = p.cin_blk * 2, rby = p.co_blk * 2;
  const int x_total = p.n_chunks * p.x_bytes;
  const int dy_box = (p.ky_b ? p.HHy : kWhRows) * kWhCols * rby;
  const int mt0 = blockIdx.x * p.mt;
  const int my_mt = min(p.mt, p.n_tiles_m - mt0);
  const int co0 = p.ky_b ? 0 : blockIdx.y * p.n_cols;
  // ky_b: highest filter row of this CTA's group (atom 0); rows ky_hi - j, j = 0 .. g_ky-1 (negative = padding)
  const int ky_hi = p.ky_b ? min(p.kh, (int)(blockIdx.y + 1) * p.g_ky) - 1 : 0;
  const int total_blocks = p.n_img * p.blocks_w * p.blocks_h;
  const int t_beg = blockIdx.z * p.blocks_per_split;
  const int t_end = min(total_blocks, t_beg + p.blocks_per_split);
  uint32_t tmem_cols = 32;
  while ((int)tmem_cols < p.mt * p.n_cols) tmem_cols <<= 1;

  if (threadIdx.x == 0) {
    prefetch_tmap(&map_x);
    prefetch_tmap(&map_dy);
    for (int s = 0; s < p.stages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(&tmem_full_bar, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(&tmem_base_s, tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_s;

  if (warp == 0) {
    // ===== TMA producer: x halo (all channel chunks) + dy block per stage =====
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      const uint32_t tx_bytes = (uint32_t)(p.n_chunks * p.HHx * p.HWx * rbx + p.b_boxes * dy_box);
      for (int t = t_beg; t < t_end; ++t) {
        const int n = t / (p.blocks_w * p.blocks_h);
        const int r = t - n * (p.blocks_w * p.blocks_h);
        const int oy0 = (r / p.blocks_w) * kWhRows, ox0 = (r % p.blocks_w) * kWhCols;
        mbar_wait(&empty_bar[stage], phase ^ 1);
        mbar_arrive_expect_tx(&full_bar[stage], tx_bytes);
        uint8_t* sx = smem + stage * p.stage_bytes;
        for (int c = 0; c < p.n_chunks; ++c)
          tma_load_4d(sx + c * p.x_bytes, &map_x, &full_bar[stage], c * p.cin_blk, ox0 + p.min_dw, oy0 + p.min_dh, n);
        uint8_t* sy = sx + x_total;
        if (p.ky_b) {
          // block rows are INPUT rows iy = oy0 + min_dh ..; dy row of atom j at block row y: iy + pad - (ky_hi - j)
          tma_load_4d(sy, &map_dy, &full_bar[stage], 0, ox0, oy0 + p.min_dh + p.pad - ky_hi, n);
        } else {
          for (int b = 0; b < p.b_boxes; ++b)
            tma_load_4d(sy + b * dy_box, &map_dy, &full_bar[stage], co0 + b * p.co_blk, ox0, oy0, n);
        }
        if (++stage == p.stages) { stage = 0; phase ^= 1; }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ===== MMA issuer: one elected lane waits and issues =====
    const uint32_t idesc = umma_idesc_bf16(128, p.n_cols, 1, 1);
    const uint32_t a_hi = desc_hi(8 * rbx, rbx), b_hi = desc_hi(8 * rby, rby);
    const uint32_t a_ystep = (uint32_t)(p.HWx * rbx) >> 4;          // next block row: one halo line down
    const uint32_t b_ystep = (uint32_t)(kWhCols * rby) >> 4;        // next block row: 16 dy rows down
    const uint32_t b_lbo = ((uint32_t)(p.ky_b ? kWhCols * rby : dy_box) >> 4) << 16;
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0, first = 0;
      for (int t = t_beg; t < t_end; ++t) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        const uint32_t sx = smem_u32(smem + stage * p.stage_bytes);
        const uint32_t x_lo = (sx >> 4) & 0x3FFF;
        const uint32_t y_lo = (((sx + x_total) >> 4) & 0x3FFF) | b_lbo;
        for (int m = 0; m < my_mt; ++m) {
          const uint32_t a_lo = x_lo + p.tiles[mt0 + m].lo_const;
          const uint32_t acc = tmem_base + m * p.n_cols;
#pragma unroll
          for (int y = 0; y < kWhRows; ++y)
            umma_bf16_split(acc, a_lo + y * a_ystep, a_hi, y_lo + y * b_ystep, b_hi, idesc, first | (uint32_t)y);
        }
        umma_commit(&empty_bar[stage]);
        first = 1;
        if (++stage == p.stages) { stage = 0; phase ^= 1; }
      }
      umma_commit(&tmem_full_bar);
    }
    __syncwarp();
  } else if (warp >= 2 && t_end > t_beg) {
    // ===== epilogue: TMEM -> red.global.add.f32x4 =====
    const int q = warp & 3;
    const int row = q * 32 + lane;                  // accumulator lane
    const int atom = row / p.cin_blk, cin = row % p.cin_blk;
    mbar_wait(&tmem_full_bar, 0);
    tc_fence_after();
    for (int m = 0; m < my_mt; ++m) {
      const int tap = p.tiles[mt0 + m].tap[atom];      // ky_b: the kx of this atom (filter row comes from the column)
      const int c0 = p.tiles[mt0 + m].c0[atom];
      float* dst0 = (tap >= 0) ? p.dw + ((int64_t)tap * p.Ci + c0 + cin) * p.Co + co0 : nullptr;
      for (int c = 0; c < p.n_cols; c += 16) {
        uint32_t v[16];
        tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(m * p.n_cols + c), v);
        tmem_ld_wait();
        float* dst = dst0;
        if (p.ky_b) {
          const int ky = ky_hi - c / p.Co;               // 16-column chunks never straddle a filter row (Co % 16 == 0)
          const bool mine = ky >= 0 && ky > ky_hi - p.g_ky && ky >= (int)blockIdx.y * p.g_ky;
          dst = (tap >= 0 && mine) ? p.dw + ((int64_t)(ky * p.kw + tap) * p.Ci + c0 + cin) * p.Co + (c % p.Co) - c : nullptr;
        }
        if (dst != nullptr) {
#pragma unroll
          for (int j = 0; j < 16; j += 4) {
            const float4 f = make_float4(__uint_as_float(v[j]), __uint_as_float(v[j + 1]), __uint_as_float(v[j + 2]),
                                         __uint_as_float(v[j + 3]));
            atomicAdd(reinterpret_cast<float4*>(dst + c + j), f);
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, tmem_cols);
}

bool conv_wgrad_halo_supported(const p2p_conv* c) {
  const bool ci_ok = c->Ci == 16 || c->Ci == 32 || c->Ci == 64 || c->Ci == 128;
  const bool co_ok = c->Co == 16 || c->Co == 32 || c->Co == 64 || c->Co == 128 || c->Co == 256;
  return c->stride == 1 && ci_ok && co_ok && c->kh <= 9 && c->kw <= 9;
}

int conv_wgrad_halo(const p2p_conv* c, const void* in, const void* dout, float* dw, cudaStream_t st) {
  WhParams p{};
  const int cin_blk = c->Ci >= 64 ? 64 : c->Ci;
  const int rbx = cin_blk * 2;
  const int n_chunks = c->Ci >= 64 ? c->Ci / 64 : 1;
  const int apt = 128 / cin_blk;                      // atoms per M tile
  int extra_w = 0, extra_h = 0;                       // halo slack read by padding atoms
  int nt = 0;
  auto new_tile = [&]() -> WhTile& {
    WhTile& t = p.tiles[nt++];
    for (int j = 0; j < 8; ++j) { t.tap[j] = -1; t.c0[j] = 0; }
    return t;
  };
  static const bool no_kyb = getenv("P2P_NO_WGRAD_KYB") != nullptr && atoi(getenv("P2P_NO_WGRAD_KYB")) != 0;
  const bool ky_b = !no_kyb && c->Ci <= 64 && c->Co <= 64 && c->kh > 1;
  if (ky_b) {
    // M tiles chain the kx taps only; filter rows ride on the N dimension (see WhParams)
    for (int kx = 0; kx < c->kw; kx += apt) {
      if (nt >= kWhMaxTiles) return -2;
      WhTile& t = new_tile();
      t.lo_const = (uint32_t)kx | 0x40000000u;                             // kx chain starting at column kx, row 0
      for (int j = 0; j < apt && kx + j < c->kw; ++j) t.tap[j] = (int16_t)(kx + j);
      if (kx + apt > c->kw) extra_w = kx + apt - c->kw;
    }
  } else if (c->Ci == 128) {
    // one tile per tap: atoms = the two 64-channel halves, one halo buffer apart (filled in below)
    for (int ky = 0; ky < c->kh; ++ky)
      for (int kx = 0; kx < c->kw; ++kx) {
        if (nt >= kWhMaxTiles) return -2;
        WhTile& t = new_tile();
        t.lo_const = (uint32_t)(ky * 1000 + kx);      // placeholder: (ky, kx), resolved once HWx is known
        t.tap[0] = t.tap[1] = (int16_t)(ky * c->kw + kx);
        t.c0[0] = 0; t.c0[1] = 64;
      }
  } else {
    // chains of `apt` consecutive taps along kx; the leftover columns are chained along ky
    const int full = (c->kw / apt) * apt;
    for (int ky = 0; ky < c->kh; ++ky)
      for (int kx = 0; kx < full; kx += apt) {
        if (nt >= kWhMaxTiles) return -2;
        WhTile& t = new_tile();
        t.lo_const = (uint32_t)(ky * 1000 + kx) | 0x40000000u;            // kx chain
        for (int j = 0; j < apt; ++j) t.tap[j] = (int16_t)(ky * c->kw + kx + j);
      }
    for (int kx = full; kx < c->kw; ++kx)
      for (int ky = 0; ky < c->kh; ky += apt) {
        if (nt >= kWhMaxTiles) return -2;
        WhTile& t = new_tile();
        t.lo_const = (uint32_t)(ky * 1000 + kx) | 0x80000000u;            // ky chain
        for (int j = 0; j < apt && ky + j < c->kh; ++j) t.tap[j] = (int16_t)((ky + j) * c->kw + kx);
        if (ky + apt > c->kh) extra_h = extra_h > ky + apt - c->kh ? extra_h : ky + apt - c->kh;
      }
  }
  p.n_tiles_m = nt;
  p.cin_blk = cin_blk;
  p.n_chunks = n_chunks;
  p.min_dh = -c->pad;
  p.min_dw = -c->pad;
  p.HWx = kWhCols + c->kw - 1 + extra_w;
  p.HHx = ky_b ? kWhRows : kWhRows + c->kh - 1 + extra_h;
  p.ky_b = ky_b ? 1 : 0;
  p.kh = c->kh; p.kw = c->kw; p.pad = c->pad;
  int ky_groups = 1;
  if (ky_b) {
    const int max_rows = 256 / c->Co;                                     // filter rows that fit one N <= 256
    ky_groups = (c->kh + max_rows - 1) / max_rows;
    p.g_ky = (c->kh + ky_groups - 1) / ky_groups;
    p.HHy = kWhRows + p.g_ky - 1;
  }
  p.x_bytes = ((p.HHx * p.HWx * rbx + 1023) / 1024) * 1024;
  for (int i = 0; i < nt; ++i) {
    const uint32_t code = p.tiles[i].lo_const;
    const int ky = (int)(code & 0x3FFFFFFFu) / 1000, kx = (int)(code & 0x3FFFFFFFu) % 1000;
    uint32_t lbo_bytes;
    if (c->Ci == 128) lbo_bytes = (uint32_t)p.x_bytes;
    else if (code & 0x40000000u) lbo_bytes = (uint32_t)rbx;
    else lbo_bytes = (uint32_t)(p.HWx * rbx);
    p.tiles[i].lo_const = ((lbo_bytes >> 4) << 16) | (uint32_t)(((ky * p.HWx + kx) * rbx) >> 4);
    if ((lbo_bytes >> 4) > 0x3FFF) return -2;
  }
  p.co_blk = c->Co >= 64 ? 64 : c->Co;
  p.n_cols = ky_b ? p.g_ky * c->Co : (c->Co >= 256 ? 256 : c->Co);
  p.b_boxes = ky_b ? 1 : p.n_cols / p.co_blk;
  int mt_max = 512 / p.n_cols;
  const int groups = (nt + mt_max - 1) / mt_max;
  p.mt = (nt + groups - 1) / groups;
  p.stage_bytes = n_chunks * p.x_bytes +
                  ((p.b_boxes * (ky_b ? p.HHy : kWhRows) * kWhCols * p.co_blk * 2 + 1023) / 1024) * 1024;
  int stages = (216 * 1024) / p.stage_bytes;
  stages = stages > 6 ? 6 : stages;
  if (stages < 2) return -2;
  p.stages = stages;
  p.blocks_w = (c->Wo + kWhCols - 1) / kWhCols;
  p.blocks_h = ((ky_b ? c->Ho + c->kh - 1 : c->Ho) + kWhRows - 1) / kWhRows;   // ky_b: blocks of INPUT rows
  p.n_img = c->N;
  p.Ci = c->Ci;
  p.Co = c->Co;
  p.dw = dw;
  const int total_blocks = c->N * p.blocks_w * p.blocks_h;
  const int n_tiles_n = ky_b ? ky_groups : c->Co / p.n_cols;
  int splits = kSMs / (groups * n_tiles_n);
  if (splits < 1) splits = 1;
  if (splits > total_blocks) splits = total_blocks;
  p.blocks_per_split = (total_blocks + splits - 1) / splits;
  splits = (total_blocks + p.blocks_per_split - 1) / p.blocks_per_split;

  CUtensorMap mx, my;
  {
    const uint64_t dims[4] = {(uint64_t)c->Ci, (uint64_t)c->Wi, (uint64_t)c->Hi, (uint64_t)c->N};
    const uint64_t strides[3] = {(uint64_t)c->Ci * 2, (uint64_t)c->Wi * c->Ci * 2, (uint64_t)c->Hi * c->Wi * c->Ci * 2};
    const uint32_t box[4] = {(uint32_t)cin_blk, (uint32_t)p.HWx, (uint32_t)p.HHx, 1};
    const uint32_t es[4] = {1, 1, 1, 1};
    if (int e = make_map_bf16(&mx, in, 4, dims, strides, box, es, rbx)) return e;
  }
  {
    const uint64_t dims[4] = {(uint64_t)c->Co, (uint64_t)c->Wo, (uint64_t)c->Ho, (uint64_t)c->N};
    const uint64_t strides[3] = {(uint64_t)c->Co * 2, (uint64_t)c->Wo * c->Co * 2, (uint64_t)c->Ho * c->Wo * c->Co * 2};
    const uint32_t box[4] = {(uint32_t)p.co_blk, (uint32_t)kWhCols, (uint32_t)(ky_b ? p.HHy : kWhRows), 1};
    const uint32_t es[4] = {1, 1, 1, 1};
    if (int e = make_map_bf16(&my, dout, 4, dims, strides, box, es, p.co_blk * 2)) return e;
  }
  const int smem = stages * p.stage_bytes + 1024;
  static SmemOptIn optin;
  P2P_CUDA(ensure_dyn_smem(conv_wgrad_halo_kernel, optin, smem));
  trace_kernel("conv_wgrad_halo cin_blk=%d chunks=%d ky_b=%d n_cols=%d mt=%d stages=%d", cin_blk, n_chunks, p.ky_b,
               p.n_cols, p.mt, stages);
  conv_wgrad_halo_kernel<<<dim3(groups, n_tiles_n, splits), 192, smem, st>>>(mx, my, p);
  P2P_LAUNCH_CHECK();
}

}  // namespace p2p
