// tcgen05 implicit-GEMM convolution, forward and data-gradient (sm_100a).
//
// GEMM view (reference layers: networks.py:404 ConvLayer, :422 UpsampleConvLayer, :766-784 D convs):
//   D[128 pixels x BLOCK_N channels] = sum over (tap, channel chunk) A[128 x CIN_BLK] * B[BLOCK_N x CIN_BLK]^T
// * A: one TMA 4-D box {CIN_BLK, tw, th, 1} of the NHWC activation tensor per (tap, chunk); the 128 rows are
//   a th x tw rectangle of output pixels, the tap only shifts the box origin, the conv stride is the
//   tensor map's element stride, zero padding is TMA out-of-bounds fill.  Reflection / upsample halos are
//   real data written by the producer layer (p2p_bn_act_store), so G layers need no special casing here.
// * B: TMA 2-D box of the packed weight matrix [channels_out][taps * channels_in] (K contiguous).
// * both land in shared memory in the canonical K-major swizzled layout, tcgen05.mma (M=128, N=BLOCK_N,
//   K=16) accumulates in TMEM; a 3..8 stage mbarrier ring couples one TMA thread to one MMA thread.
// * epilogue (4 warps): tcgen05.ld -> +bias -> LeakyReLU -> bf16 -> staged in the (now idle) pipeline
//   smem -> per-channel sum / sum-of-squares for train-mode BatchNorm -> coalesced 16-byte global stores.
// The data gradient is the same kernel: A = dout, B = the [ci][tap][co] packing, taps negated; a stride-2
// layer is done as 4 output-parity phases, each a stride-1 problem writing every other row / column.
#include "tc_common.cuh"

namespace p2p {

using namespace tc;

constexpr int kMaxTaps = 81;

struct TcTap {
  int16_t dh, dw;
  int32_t wk;  // first K index (element) of this tap in the weight matrix
};

struct TcConvParams {
  int tiles_w, tiles_h, tw_log2, tw;  // tile rectangle th x tw = 128 output positions
  int in_mul;                         // input coordinate = output coordinate * in_mul + tap offset
  int n_taps, c_chunks;
  int Hv, Wv;                         // valid extent of the output-position grid
  int64_t out_sN, out_sH, out_sW;     // output strides in elements
  int h_mul, h_off, w_mul, w_off;     // output row = pos * h_mul + h_off
  int co_real, act;
  __nv_bfloat16* out;
  const float* bias;
  double* stats;  // [2][Cn] or null
  int Cn;         // stored output channels
  TcTap taps[kMaxTaps];
};

template <int BLOCK_N, int CIN_BLK>
struct TcCfg {
  static constexpr int A_BYTES = 128 * CIN_BLK * 2;
  static constexpr int B_BYTES = BLOCK_N * CIN_BLK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int RAW = (96 * 1024) / STAGE_BYTES;
  static constexpr int STAGES = (BLOCK_N == 256 && CIN_BLK == 64) ? 4 : (RAW > 8 ? 8 : (RAW < 2 ? 2 : RAW));
  static constexpr int PITCH = BLOCK_N * 2 + 16;  // epilogue staging row pitch (bytes), bank-conflict free
  static constexpr int SMEM = STAGES * STAGE_BYTES + 1024;
  static constexpr int TMEM_COLS = BLOCK_N < 32 ? 32 : BLOCK_N;
  static_assert(128 * PITCH + 8192 <= STAGES * STAGE_BYTES, "epilogue staging must fit in the pipeline buffers");
};

template <int BLOCK_N, int CIN_BLK>
__global__ void __launch_bounds__(192)
    conv_tc_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                   const __grid_constant__ TcConvParams p) {
  using Cfg = TcCfg<BLOCK_N, CIN_BLK>;
  constexpr int STAGES = Cfg::STAGES;
  constexpr int ROW_BYTES = CIN_BLK * 2;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t full_bar[STAGES], empty_bar[STAGES], tmem_full_bar;
  __shared__ uint32_t tmem_base_s;
  __shared__ __align__(16) float s_bias[BLOCK_N];

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tiles_per_img = p.tiles_w * p.tiles_h;
  const int n_img = blockIdx.x / tiles_per_img;
  const int trem = blockIdx.x - n_img * tiles_per_img;
  const int tyi = trem / p.tiles_w, txi = trem - tyi * p.tiles_w;
  const int th = 128 >> p.tw_log2;
  const int oy0 = tyi * th, ox0 = txi * p.tw;
  const int n0 = blockIdx.y * BLOCK_N;
  const int num_kb = p.n_taps * p.c_chunks;

  if (threadIdx.x == 0) {
    prefetch_tmap(&map_a);
    prefetch_tmap(&map_b);
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(&tmem_full_bar, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(&tmem_base_s, Cfg::TMEM_COLS);
  if (warp >= 2) {
    for (int i = threadIdx.x - 64; i < BLOCK_N; i += 128)
      s_bias[i] = (p.bias != nullptr && n0 + i < p.co_real) ? p.bias[n0 + i] : 0.f;
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_s;

  if (warp == 0) {
    // ===== TMA producer (warp-uniform loop, one elected lane issues: keeps TMA operands in uniform registers) =====
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int kb = 0; kb < num_kb; ++kb) {
        const int tap = kb / p.c_chunks, cc = kb - tap * p.c_chunks;
        mbar_wait(&empty_bar[stage], phase ^ 1);
        mbar_arrive_expect_tx(&full_bar[stage], Cfg::STAGE_BYTES);
        uint8_t* sa = smem + stage * Cfg::STAGE_BYTES;
        tma_load_4d(sa, &map_a, &full_bar[stage], cc * CIN_BLK, ox0 * p.in_mul + p.taps[tap].dw,
                    oy0 * p.in_mul + p.taps[tap].dh, n_img);
        tma_load_2d(sa + Cfg::A_BYTES, &map_b, &full_bar[stage], p.taps[tap].wk + cc * CIN_BLK, n0);
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ===== MMA issuer =====
    constexpr uint32_t idesc = umma_idesc_bf16(128, BLOCK_N, 0, 0);
    const uint32_t d_hi = desc_hi(8 * ROW_BYTES, ROW_BYTES);      // same geometry for A and B tiles
    if (elect_one()) {                                            // one lane waits and issues for the whole tile
      int stage = 0;
      uint32_t phase = 0;
      for (int kb = 0; kb < num_kb; ++kb) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        const uint32_t sa = smem_u32(smem + stage * Cfg::STAGE_BYTES);
        const uint32_t a_lo = desc_lo(sa, 16), b_lo = desc_lo(sa + Cfg::A_BYTES, 16);
#pragma unroll
        for (int k = 0; k < CIN_BLK / 16; ++k)
          umma_bf16_split(tmem_base, a_lo + 2 * k, d_hi, b_lo + 2 * k, d_hi, idesc, (uint32_t)(kb | k));
        umma_commit(&empty_bar[stage]);  // frees this stage once the MMAs above have read it
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
      umma_commit(&tmem_full_bar);
    }
    __syncwarp();
  } else if (warp >= 2) {
    // ===== epilogue =====
    mbar_wait(&tmem_full_bar, 0);
    tc_fence_after();
    // pipeline buffers are idle now (every MMA that read them has completed): reuse them as staging
    uint8_t* stg = smem;
    float* s_red = reinterpret_cast<float*>(smem + 128 * Cfg::PITCH);
    EpiOut e;
    e.out = p.out + (int64_t)n_img * p.out_sN;
    e.out_sH = p.out_sH; e.out_sW = p.out_sW;
    e.h_mul = p.h_mul; e.h_off = p.h_off; e.w_mul = p.w_mul; e.w_off = p.w_off;
    e.Hv = p.Hv; e.Wv = p.Wv;
    e.oy0 = oy0; e.ox0 = ox0; e.tw_log2 = p.tw_log2;
    e.n0 = n0; e.co_real = p.co_real; e.Cn = p.Cn; e.act = p.act;
    e.stats = p.stats;
    epilogue_tile<BLOCK_N>(tmem_base, warp & 3, lane, threadIdx.x - 64, stg, s_bias, s_red, e, nullptr);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
}

// ---------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------
static bool pow2_16_512(int c) { return c == 16 || c == 32 || c == 64 || c == 128 || c == 256 || c == 512; }

bool conv_fwd_tc_supported(const p2p_conv* c) {
  return pow2_16_512(c->Ci) && pow2_16_512(c->Co) && c->kh * c->kw <= kMaxTaps && (c->stride == 1 || c->stride == 2);
}
bool conv_dgrad_tc_supported(const p2p_conv* c) {
  return conv_fwd_tc_supported(c) && c->kh >= c->stride && c->kw >= c->stride;
}

static void pick_tile(int Hv, int Wv, int max_tw, int* tw_out) {
  long best = -1;
  int best_tw = 8;
  for (int tw = 128; tw >= 8; tw >>= 1) {
    if (tw > max_tw) continue;
    const int th = 128 / tw;
    const long tiles = (long)((Wv + tw - 1) / tw) * ((Hv + th - 1) / th);
    if (best < 0 || tiles < best) { best = tiles; best_tw = tw; }
  }
  *tw_out = best_tw;
}

template <int BLOCK_N, int CIN_BLK>
static int launch_cfg(const CUtensorMap& ma, const CUtensorMap& mb, const TcConvParams& p, int grid_x, int grid_y,
                      cudaStream_t st) {
  using Cfg = TcCfg<BLOCK_N, CIN_BLK>;
  static SmemOptIn optin;
  P2P_CUDA(ensure_dyn_smem(conv_tc_kernel<BLOCK_N, CIN_BLK>, optin, Cfg::SMEM));
  trace_kernel("conv_tc<%d,%d> tw=%d stride=%d taps=%d chunks=%d", BLOCK_N, CIN_BLK, p.tw, p.in_mul, p.n_taps, p.c_chunks);
  conv_tc_kernel<BLOCK_N, CIN_BLK><<<dim3(grid_x, grid_y), 192, Cfg::SMEM, st>>>(ma, mb, p);
  P2P_LAUNCH_CHECK();
}

template <int CIN_BLK>
static int launch_n(int block_n, const CUtensorMap& ma, const CUtensorMap& mb, const TcConvParams& p, int gx, int gy,
                    cudaStream_t st) {
  switch (block_n) {
    case 16: return launch_cfg<16, CIN_BLK>(ma, mb, p, gx, gy, st);
    case 32: return launch_cfg<32, CIN_BLK>(ma, mb, p, gx, gy, st);
    case 64: return launch_cfg<64, CIN_BLK>(ma, mb, p, gx, gy, st);
    case 128: return launch_cfg<128, CIN_BLK>(ma, mb, p, gx, gy, st);
    default: return launch_cfg<256, CIN_BLK>(ma, mb, p, gx, gy, st);
  }
}

// One implicit-GEMM problem: A source tensor [N,Hs,Ws,Ck], B matrix [Cn][ktot], output positions Hv x Wv.
// conv_halo.cu
int run_halo_problem(const void* a_src, int N, int Hs, int Ws, int Ck, const void* b_mat, int Cn, int64_t ktot,
                     const int* dh, const int* dw, const int* wk, int n_taps, int Hv, int Wv, __nv_bfloat16* out,
                     int64_t out_sN, int64_t out_sH, int64_t out_sW, int h_mul, int h_off, int w_mul, int w_off,
                     int co_real, int act, const float* bias, double* stats, bool force, cudaStream_t st);

static bool g_use_halo = true;    // P2P_NO_HALO=1 forces the per-tap kernel, P2P_FORCE_HALO=1 the halo kernel
static bool g_force_halo = false;  // (A/B comparisons; by default run_halo_problem picks per shape)

static int run_problem(const void* a_src, int N, int Hs, int Ws, int Ck, const void* b_mat, int Cn, int64_t ktot,
                       TcConvParams& p, cudaStream_t st) {
  static bool env_checked = false;
  if (!env_checked) {
    const char* e = getenv("P2P_NO_HALO");
    g_use_halo = !(e != nullptr && e[0] == '1');
    const char* f = getenv("P2P_FORCE_HALO");
    g_force_halo = f != nullptr && f[0] == '1';
    env_checked = true;
  }
  if (p.in_mul == 1 && g_use_halo) {
    // stride-1 problem: halo-resident kernel (one input read per channel chunk instead of one per tap)
    int dh[kMaxTaps], dw[kMaxTaps], wk[kMaxTaps];
    for (int t = 0; t < p.n_taps; ++t) { dh[t] = p.taps[t].dh; dw[t] = p.taps[t].dw; wk[t] = p.taps[t].wk; }
    const int rc = run_halo_problem(a_src, N, Hs, Ws, Ck, b_mat, Cn, ktot, dh, dw, wk, p.n_taps, p.Hv, p.Wv, p.out,
                                    p.out_sN, p.out_sH, p.out_sW, p.h_mul, p.h_off, p.w_mul, p.w_off, p.co_real, p.act,
                                    p.bias, p.stats, g_force_halo, st);
    if (rc != -2) return rc;
  }
  const int cin_blk = Ck >= 64 ? 64 : Ck;
  const int block_n = Cn >= 256 ? 256 : Cn;
  int tw;
  pick_tile(p.Hv, p.Wv, 256 / p.in_mul, &tw);
  const int th = 128 / tw;
  p.tw = tw;
  p.tw_log2 = 31 - __builtin_clz((unsigned)tw);
  p.tiles_w = (p.Wv + tw - 1) / tw;
  p.tiles_h = (p.Hv + th - 1) / th;
  p.c_chunks = Ck / cin_blk;
  p.Cn = Cn;

  CUtensorMap ma, mb;
  {
    const uint64_t dims[4] = {(uint64_t)Ck, (uint64_t)Ws, (uint64_t)Hs, (uint64_t)N};
    const uint64_t strides[3] = {(uint64_t)Ck * 2, (uint64_t)Ws * Ck * 2, (uint64_t)Hs * Ws * Ck * 2};
    const uint32_t box[4] = {(uint32_t)cin_blk, (uint32_t)(tw * p.in_mul), (uint32_t)(th * p.in_mul), 1};
    const uint32_t es[4] = {1, (uint32_t)p.in_mul, (uint32_t)p.in_mul, 1};
    if (int e = make_map_bf16(&ma, a_src, 4, dims, strides, box, es, cin_blk * 2)) return e;
  }
  {
    const uint64_t dims[2] = {(uint64_t)ktot, (uint64_t)Cn};
    const uint64_t strides[1] = {(uint64_t)ktot * 2};
    const uint32_t box[2] = {(uint32_t)cin_blk, (uint32_t)block_n};
    const uint32_t es[2] = {1, 1};
    if (int e = make_map_bf16(&mb, b_mat, 2, dims, strides, box, es, cin_blk * 2)) return e;
  }
  const int gx = p.tiles_w * p.tiles_h * N, gy = Cn / block_n;
  switch (cin_blk) {
    case 16: return launch_n<16>(block_n, ma, mb, p, gx, gy, st);
    case 32: return launch_n<32>(block_n, ma, mb, p, gx, gy, st);
    default: return launch_n<64>(block_n, ma, mb, p, gx, gy, st);
  }
}

int conv_fwd_tc(const p2p_conv* c, const void* in, const void* w, const float* bias, void* out, double* stats,
                cudaStream_t st) {
  TcConvParams p{};
  p.in_mul = c->stride;
  p.n_taps = c->kh * c->kw;
  for (int ky = 0; ky < c->kh; ++ky)
    for (int kx = 0; kx < c->kw; ++kx) {
      TcTap& t = p.taps[ky * c->kw + kx];
      t.dh = (int16_t)(ky - c->pad);
      t.dw = (int16_t)(kx - c->pad);
      t.wk = (ky * c->kw + kx) * c->Ci;
    }
  p.Hv = c->Ho;
  p.Wv = c->Wo;
  p.out_sN = (int64_t)c->Ho * c->Wo * c->Co;
  p.out_sH = (int64_t)c->Wo * c->Co;
  p.out_sW = c->Co;
  p.h_mul = p.w_mul = 1;
  p.h_off = p.w_off = 0;
  p.co_real = c->co_real;
  p.act = c->act;
  p.out = (__nv_bfloat16*)out;
  p.bias = bias;
  p.stats = stats;
  return run_problem(in, c->N, c->Hi, c->Wi, c->Ci, w, c->Co, (int64_t)c->kh * c->kw * c->Ci, p, st);
}

int conv_dgrad_tc(const p2p_conv* c, const void* dout, const void* wt, void* din, cudaStream_t st) {
  const int s = c->stride;
  for (int py = 0; py < s; ++py)
    for (int px = 0; px < s; ++px) {
      TcConvParams p{};
      p.in_mul = 1;
      int nt = 0;
      for (int ky = 0; ky < c->kh; ++ky) {
        if ((py + c->pad - ky) % s != 0) continue;
        for (int kx = 0; kx < c->kw; ++kx) {
          if ((px + c->pad - kx) % s != 0) continue;
          TcTap& t = p.taps[nt++];
          t.dh = (int16_t)((py + c->pad - ky) / s);
          t.dw = (int16_t)((px + c->pad - kx) / s);
          t.wk = (ky * c->kw + kx) * c->Co;
        }
      }
      P2P_REQUIRE(nt > 0, "a stride phase without taps");
      p.n_taps = nt;
      p.Hv = (c->Hi - py + s - 1) / s;
      p.Wv = (c->Wi - px + s - 1) / s;
      if (p.Hv <= 0 || p.Wv <= 0) continue;
      p.out_sN = (int64_t)c->Hi * c->Wi * c->Ci;
      p.out_sH = (int64_t)c->Wi * c->Ci;
      p.out_sW = c->Ci;
      p.h_mul = p.w_mul = s;
      p.h_off = py;
      p.w_off = px;
      p.co_real = c->Ci;  // every stored input channel gets its gradient (pad channels come out 0: zero weights)
      p.act = P2P_ACT_NONE;
      p.out = (__nv_bfloat16*)din;
      p.bias = nullptr;
      p.stats = nullptr;
      if (int e = run_problem(dout, c->N, c->Ho, c->Wo, c->Co, wt, c->Ci, (int64_t)c->kh * c->kw * c->Co, p, st)) return e;
    }
  return 0;
}

}  // namespace p2p
