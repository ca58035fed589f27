// tcgen05 weight-gradient kernel (sm_100a):  dW[tap][ci][co] += sum over output pixels x[p+tap][ci] * dy[p][co]
// (backward of nn.Conv2d w.r.t. its weight: reference networks.py:400,416,766-784 through autograd).
//
// GEMM view: D[M = (tap, ci) rows][N = co] with the reduction over PIXELS.  In NHWC the pixel index is the
// slow index of both operands, so both are "MN-major" UMMA operands: a TMA box {channels, tw, th, 1} lands
// in shared memory as PIX rows of one swizzle span each -- exactly the canonical MN-major layout
// (rows = K, 8-row groups SBO apart, the next block of channels / the next tap one box (LBO) further on).
//   * one M tile (128 TMEM lanes) = 128/CIN_BLK boxes: two 64-channel chunks of one tap, or 2/4/8 taps of a
//     64/32/16-channel layer; a CTA owns MT such tiles (MT x N fp32 accumulator columns <= 512) that all
//     reuse the same dy tile, and one slice of the pixel range (split-K);
//   * TMA producer thread -> mbarrier ring -> MMA thread (tcgen05.mma, M=128, N=Co tile, K=16 pixels)
//     -> tcgen05.commit; the 4 epilogue warps read TMEM and red.global.add (float4) into dW.
#include "tc_common.cuh"

namespace p2p {

using namespace tc;

constexpr int kMaxSlots = 160;

struct WgSlot {      // one TMA box of the A operand = 128/boxes_per_tile TMEM lanes
  int8_t dh, dw;     // tap offset (input coordinate = out * stride + d - pad folded in)
  int16_t c0;        // first input channel of the box
  int16_t tap;       // flattened tap index ky*kw+kx, or -1 for a padding slot whose result is discarded
  int16_t pad_;
};

struct WgParams {
  int tiles_w, tiles_h, tw, th, n_img;   // pixel tiles: th x tw = PIX output pixels
  int pix;                               // 32 or 64
  int in_mul;                            // conv stride
  int mt;                                // M tiles per CTA
  int m_tiles_total;
  int boxes_per_tile;                    // 128 / cin_blk
  int cin_blk, co_blk, n_cols;           // channels per A box, channels per B box, UMMA N
  int b_boxes;                           // n_cols / co_blk
  int stages;
  int Ci, Co;
  int tiles_per_split;
  float* dw;
  WgSlot slots[kMaxSlots];
};

__global__ void __launch_bounds__(192)
    conv_wgrad_tc_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_dy,
                         const __grid_constant__ WgParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t full_bar[8], empty_bar[8], tmem_full_bar;
  __shared__ uint32_t tmem_base_s;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int a_row_bytes = p.cin_blk * 2, b_row_bytes = p.co_blk * 2;
  const int a_box_bytes = p.pix * a_row_bytes, b_box_bytes = p.pix * b_row_bytes;
  const int b_bytes = p.b_boxes * b_box_bytes;             // dy tile
  const int a_tile_bytes = p.boxes_per_tile * a_box_bytes;  // one M tile (always pix * 256 B)
  const int mt0 = blockIdx.x * p.mt;
  const int my_mt = min(p.mt, p.m_tiles_total - mt0);
  const int stage_bytes = b_bytes + p.mt * a_tile_bytes;
  const int co0 = blockIdx.y * p.n_cols;
  const int total_tiles = p.n_img * p.tiles_w * p.tiles_h;
  const int t_beg = blockIdx.z * p.tiles_per_split;
  const int t_end = min(total_tiles, t_beg + p.tiles_per_split);
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
    // ===== TMA producer =====
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      const uint32_t tx_bytes = (uint32_t)(b_bytes + my_mt * a_tile_bytes);
      for (int t = t_beg; t < t_end; ++t) {
        const int n = t / (p.tiles_w * p.tiles_h);
        const int r = t - n * (p.tiles_w * p.tiles_h);
        const int oy0 = (r / p.tiles_w) * p.th, ox0 = (r % p.tiles_w) * p.tw;
        mbar_wait(&empty_bar[stage], phase ^ 1);
        mbar_arrive_expect_tx(&full_bar[stage], tx_bytes);
        uint8_t* sb = smem + stage * stage_bytes;
        for (int b = 0; b < p.b_boxes; ++b)
          tma_load_4d(sb + b * b_box_bytes, &map_dy, &full_bar[stage], co0 + b * p.co_blk, ox0, oy0, n);
        uint8_t* sa = sb + b_bytes;
        for (int m = 0; m < my_mt; ++m)
          for (int b = 0; b < p.boxes_per_tile; ++b) {
            const WgSlot s = p.slots[(mt0 + m) * p.boxes_per_tile + b];
            tma_load_4d(sa + m * a_tile_bytes + b * a_box_bytes, &map_x, &full_bar[stage], s.c0,
                        ox0 * p.in_mul + s.dw, oy0 * p.in_mul + s.dh, n);
          }
        if (++stage == p.stages) { stage = 0; phase ^= 1; }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ===== MMA issuer: one elected lane waits and issues =====
    const uint32_t idesc = umma_idesc_bf16(128, p.n_cols, 1, 1);
    const uint32_t a_hi = desc_hi(8 * a_row_bytes, a_row_bytes), b_hi = desc_hi(8 * b_row_bytes, b_row_bytes);
    const uint32_t a_kstep = (uint32_t)(16 * a_row_bytes) >> 4, b_kstep = (uint32_t)(16 * b_row_bytes) >> 4;
    const uint32_t a_tstep = (uint32_t)a_tile_bytes >> 4;
    const int ksteps = p.pix / 16;
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      uint32_t first = 0;
      for (int t = t_beg; t < t_end; ++t) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        const uint32_t sb = smem_u32(smem + stage * stage_bytes);
        const uint32_t b_lo0 = desc_lo(sb, (uint32_t)b_box_bytes);
        uint32_t a_lo = desc_lo(sb + b_bytes, (uint32_t)a_box_bytes);
        for (int m = 0; m < my_mt; ++m, a_lo += a_tstep) {
          const uint32_t acc = tmem_base + m * p.n_cols;
          for (int ks = 0; ks < ksteps; ++ks)
            umma_bf16_split(acc, a_lo + ks * a_kstep, a_hi, b_lo0 + ks * b_kstep, b_hi, idesc, first | (uint32_t)ks);
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
    const int row = q * 32 + lane;                  // TMEM lane = row of the M tile
    const int box = row / p.cin_blk, cin = row % p.cin_blk;
    mbar_wait(&tmem_full_bar, 0);
    tc_fence_after();
    for (int m = 0; m < my_mt; ++m) {
      const WgSlot s = p.slots[(mt0 + m) * p.boxes_per_tile + box];
      float* dst = (s.tap >= 0) ? p.dw + ((int64_t)s.tap * p.Ci + s.c0 + cin) * p.Co + co0 : nullptr;
      for (int c = 0; c < p.n_cols; c += 16) {
        uint32_t v[16];
        tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(m * p.n_cols + c), v);
        tmem_ld_wait();
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

static bool pow2_16_512(int c) { return c == 16 || c == 32 || c == 64 || c == 128 || c == 256 || c == 512; }

bool conv_wgrad_tc_supported(const p2p_conv* c) {
  if (!(pow2_16_512(c->Ci) && pow2_16_512(c->Co) && (c->stride == 1 || c->stride == 2))) return false;
  const int cin_blk = c->Ci >= 64 ? 64 : c->Ci;
  const int bpt = 128 / cin_blk;
  const int units = c->kh * c->kw * (c->Ci / cin_blk);   // boxes needed
  const int m_tiles = (units + bpt - 1) / bpt;
  return m_tiles * bpt <= kMaxSlots && c->kh <= 127 && c->kw <= 127;
}

// conv_wgrad_halo.cu
bool conv_wgrad_halo_supported(const p2p_conv* c);
int conv_wgrad_halo(const p2p_conv* c, const void* in, const void* dout, float* dw, cudaStream_t st);

int conv_wgrad_tc(const p2p_conv* c, const void* in, const void* dout, float* dw, cudaStream_t st) {
  static int no_halo = -1;
  if (no_halo < 0) {
    const char* e = getenv("P2P_NO_WGRAD_HALO");
    no_halo = (e != nullptr && e[0] == '1') ? 1 : 0;
  }
  if (!no_halo && conv_wgrad_halo_supported(c)) {
    const int rc = conv_wgrad_halo(c, in, dout, dw, st);   // stride-1 layers: one halo load per stage for all taps
    if (rc != -2) return rc;
  }
  WgParams p{};
  const int cin_blk = c->Ci >= 64 ? 64 : c->Ci;
  const int co_blk = c->Co >= 64 ? 64 : c->Co;
  const int n_cols = c->Co >= 256 ? 256 : c->Co;
  const int bpt = 128 / cin_blk;
  const int chunks = c->Ci / cin_blk;
  // A-operand boxes in (tap, chunk) order; tiles are runs of `bpt` boxes
  int ns = 0;
  for (int ky = 0; ky < c->kh; ++ky)
    for (int kx = 0; kx < c->kw; ++kx)
      for (int ch = 0; ch < chunks; ++ch) {
        WgSlot& s = p.slots[ns++];
        s.dh = (int8_t)(ky - c->pad);
        s.dw = (int8_t)(kx - c->pad);
        s.c0 = (int16_t)(ch * cin_blk);
        s.tap = (int16_t)(ky * c->kw + kx);
      }
  const int m_tiles = (ns + bpt - 1) / bpt;
  for (; ns < m_tiles * bpt; ++ns) {   // padding slots: re-read the first box, discard the result
    p.slots[ns] = p.slots[0];
    p.slots[ns].tap = -1;
  }
  int mt_max = 512 / n_cols;
  if (mt_max > 6) mt_max = 6;
  const int groups = (m_tiles + mt_max - 1) / mt_max;
  const int mt = (m_tiles + groups - 1) / groups;
  p.mt = mt;
  p.m_tiles_total = m_tiles;
  p.boxes_per_tile = bpt;
  p.cin_blk = cin_blk;
  p.co_blk = co_blk;
  p.n_cols = n_cols;
  p.b_boxes = n_cols / co_blk;
  p.pix = mt <= 3 ? 64 : 32;
  p.in_mul = c->stride;
  p.Ci = c->Ci;
  p.Co = c->Co;
  p.dw = dw;
  p.n_img = c->N;
  // pixel tile rectangle: as wide as possible (<= Wo rounded up), tw * th = pix
  int tw = 8;
  long best = -1;
  for (int cand = p.pix; cand >= 8; cand >>= 1) {
    if (cand * c->stride > 256) continue;
    const int th_c = p.pix / cand;
    const long tiles = (long)((c->Wo + cand - 1) / cand) * ((c->Ho + th_c - 1) / th_c);
    if (best < 0 || tiles < best) { best = tiles; tw = cand; }
  }
  p.tw = tw;
  p.th = p.pix / tw;
  p.tiles_w = (c->Wo + p.tw - 1) / p.tw;
  p.tiles_h = (c->Ho + p.th - 1) / p.th;
  const int stage_bytes = p.pix * (n_cols * 2 + mt * 256);
  int stages = (200 * 1024) / stage_bytes;
  stages = stages > 6 ? 6 : stages;
  P2P_REQUIRE(stages >= 2, "internal: wgrad stage does not fit");
  p.stages = stages;
  const int smem = stages * stage_bytes + 1024;
  const int total_tiles = c->N * p.tiles_w * p.tiles_h;
  const int n_tiles = c->Co / n_cols;
  int splits = (2 * kSMs + groups * n_tiles - 1) / (groups * n_tiles);
  if (splits > total_tiles) splits = total_tiles;
  if (splits < 1) splits = 1;
  p.tiles_per_split = (total_tiles + splits - 1) / splits;
  splits = (total_tiles + p.tiles_per_split - 1) / p.tiles_per_split;

  CUtensorMap mx, my;
  {
    const uint64_t dims[4] = {(uint64_t)c->Ci, (uint64_t)c->Wi, (uint64_t)c->Hi, (uint64_t)c->N};
    const uint64_t strides[3] = {(uint64_t)c->Ci * 2, (uint64_t)c->Wi * c->Ci * 2, (uint64_t)c->Hi * c->Wi * c->Ci * 2};
    const uint32_t box[4] = {(uint32_t)cin_blk, (uint32_t)(p.tw * c->stride), (uint32_t)(p.th * c->stride), 1};
    const uint32_t es[4] = {1, (uint32_t)c->stride, (uint32_t)c->stride, 1};
    if (int e = make_map_bf16(&mx, in, 4, dims, strides, box, es, cin_blk * 2)) return e;
  }
  {
    const uint64_t dims[4] = {(uint64_t)c->Co, (uint64_t)c->Wo, (uint64_t)c->Ho, (uint64_t)c->N};
    const uint64_t strides[3] = {(uint64_t)c->Co * 2, (uint64_t)c->Wo * c->Co * 2, (uint64_t)c->Ho * c->Wo * c->Co * 2};
    const uint32_t box[4] = {(uint32_t)co_blk, (uint32_t)p.tw, (uint32_t)p.th, 1};
    const uint32_t es[4] = {1, 1, 1, 1};
    if (int e = make_map_bf16(&my, dout, 4, dims, strides, box, es, co_blk * 2)) return e;
  }
  static SmemOptIn optin;
  P2P_CUDA(ensure_dyn_smem(conv_wgrad_tc_kernel, optin, 220 * 1024));
  trace_kernel("conv_wgrad_tc cin_blk=%d n_cols=%d mt=%d pix=%d tw=%d stride=%d stages=%d", cin_blk, n_cols, mt, p.pix,
               p.tw, c->stride, stages);
  conv_wgrad_tc_kernel<<<dim3(groups, n_tiles, splits), 192, smem, st>>>(mx, my, p);
  P2P_LAUNCH_CHECK();
}

}  // namespace p2p
