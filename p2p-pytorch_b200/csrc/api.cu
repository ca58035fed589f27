// Library-level entry points and the convolution dispatch (tcgen05 path vs fp32-accumulate SIMT path).
#include <atomic>
#include <cstdarg>
#include <cstring>
#include <mutex>

#include "common.cuh"

namespace p2p {

static thread_local std::string g_last_error;
void set_error(const std::string& s) { g_last_error = s; }
static std::atomic<long long> g_launches{0};
void count_launches(int n) { g_launches += n; }

static std::mutex g_trace_mu;
static std::atomic<bool> g_trace_on{false};
static std::string g_trace;
bool trace_open() { return g_trace_on.load(std::memory_order_relaxed); }
void trace_kernel(const char* fmt, ...) {
  if (!trace_open()) return;
  char buf[256];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  std::lock_guard<std::mutex> lk(g_trace_mu);
  g_trace += buf;
  g_trace += '\n';
}

// conv_simt.cu
int conv_fwd_simt(const p2p_conv* c, const void* in, const void* w, const float* bias, void* out, int dtype,
                  cudaStream_t st);
int conv_dgrad_simt(const p2p_conv* c, const void* dout, const void* wt, void* din, int dtype, cudaStream_t st);
int conv_wgrad_simt(const p2p_conv* c, const void* in, const void* dout, float* dw, int dtype, cudaStream_t st);
// conv_tc.cu
bool conv_fwd_tc_supported(const p2p_conv* c);
bool conv_dgrad_tc_supported(const p2p_conv* c);
bool conv_wgrad_tc_supported(const p2p_conv* c);
int conv_fwd_tc(const p2p_conv* c, const void* in, const void* w, const float* bias, void* out, double* stats,
                cudaStream_t st);
int conv_dgrad_tc(const p2p_conv* c, const void* dout, const void* wt, void* din, cudaStream_t st);
int conv_wgrad_tc(const p2p_conv* c, const void* in, const void* dout, float* dw, cudaStream_t st);

// conv_co1.cu: weight gradient of layers with one real output channel (a memory-bound CUDA-core sweep, chosen
// under P2P_IMPL_AUTO)
bool conv_co1_supported(const p2p_conv* c);
int conv_wgrad_co1(const p2p_conv* c, const void* in, const void* dout, float* dw, int dtype, cudaStream_t st);

static int check_conv(const p2p_conv* c) {
  P2P_REQUIRE(c != nullptr, "null descriptor");
  P2P_REQUIRE(c->N > 0 && c->Hi > 0 && c->Wi > 0 && c->Ho > 0 && c->Wo > 0, "empty tensor");
  P2P_REQUIRE(c->Ci % 16 == 0 && c->Co % 16 == 0 && c->Ci > 0 && c->Co > 0, "stored channels must be multiples of 16");
  P2P_REQUIRE(c->kh > 0 && c->kw > 0 && c->stride >= 1 && c->pad >= 0, "bad kernel geometry");
  P2P_REQUIRE(c->co_real > 0 && c->co_real <= c->Co && c->ci_real > 0 && c->ci_real <= c->Ci, "bad real channel count");
  P2P_REQUIRE((c->Ho - 1) * c->stride + c->kh - c->pad <= c->Hi + c->pad &&
                  (c->Wo - 1) * c->stride + c->kw - c->pad <= c->Wi + c->pad,
              "output larger than the padded input allows");
  P2P_REQUIRE(c->act == P2P_ACT_NONE || c->act == P2P_ACT_LRELU || c->act == P2P_ACT_RELU,
              "epilogue activation must be none, relu or leaky relu");
  return 0;
}

// compute capability 10.x of the CURRENT device, cached per ordinal (0 = not looked up yet, 1 = yes, 2 = no)
static std::atomic<int> g_tc_state[kMaxDevices];
static bool device_tc() {
  int dev = 0, major = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return false;
  const bool cached = dev >= 0 && dev < kMaxDevices;
  if (cached && g_tc_state[dev].load(std::memory_order_relaxed) != 0) return g_tc_state[dev].load() == 1;
  const bool ok = cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev) == cudaSuccess && major == 10;
  if (cached) g_tc_state[dev].store(ok ? 1 : 2);
  return ok;
}

}  // namespace p2p

using namespace p2p;

extern "C" {

int p2p_version(void) { return 100; }
const char* p2p_last_error(void) { return g_last_error.c_str(); }
int p2p_device_supports_tc(void) { return device_tc() ? 1 : 0; }
long long p2p_launch_count(void) { return g_launches.load(); }

void p2p_trace_begin(void) {
  std::lock_guard<std::mutex> lk(g_trace_mu);
  g_trace.clear();
  g_trace_on.store(true);
}
int p2p_trace_end(char* buf, int cap) {
  std::lock_guard<std::mutex> lk(g_trace_mu);
  g_trace_on.store(false);
  const int need = (int)g_trace.size() + 1;
  if (buf != nullptr && cap > 0) {
    const int n = need <= cap ? need - 1 : cap - 1;
    memcpy(buf, g_trace.data(), (size_t)n);
    buf[n] = 0;
  }
  return need;
}

int p2p_conv_fwd(const p2p_conv* c, const void* in, const void* w_packed, const float* bias, void* out, double* stats,
                 int dtype, int impl, void* stream) {
  if (int e = check_conv(c)) return e;
  P2P_REQUIRE(in && w_packed && out, "null pointer");
  P2P_REQUIRE(dtype == P2P_F32 || dtype == P2P_BF16, "unknown dtype");
  cudaStream_t st = as_stream(stream);
  const bool tc_ok = dtype == P2P_BF16 && device_tc() && conv_fwd_tc_supported(c);
  P2P_REQUIRE(impl != P2P_IMPL_TC || tc_ok, "tcgen05 path does not support this shape / dtype / device");
  if (impl != P2P_IMPL_SIMT && tc_ok) return conv_fwd_tc(c, in, w_packed, bias, out, stats, st);
  if (int e = conv_fwd_simt(c, in, w_packed, bias, out, dtype, st)) return e;
  if (stats != nullptr) {
    P2P_REQUIRE(c->act == P2P_ACT_NONE, "statistics are taken before any activation");
    return p2p_channel_stats(out, stats, (int64_t)c->N * c->Ho * c->Wo, c->Co, dtype, stream);
  }
  return 0;
}

int p2p_conv_dgrad(const p2p_conv* c, const void* dout, const void* wt_packed, void* din, int dtype, int impl,
                   void* stream) {
  if (int e = check_conv(c)) return e;
  P2P_REQUIRE(dout && wt_packed && din, "null pointer");
  P2P_REQUIRE(dtype == P2P_F32 || dtype == P2P_BF16, "unknown dtype");
  cudaStream_t st = as_stream(stream);
  const bool tc_ok = dtype == P2P_BF16 && device_tc() && conv_dgrad_tc_supported(c);
  P2P_REQUIRE(impl != P2P_IMPL_TC || tc_ok, "tcgen05 path does not support this shape / dtype / device");
  if (impl != P2P_IMPL_SIMT && tc_ok) return conv_dgrad_tc(c, dout, wt_packed, din, st);
  return conv_dgrad_simt(c, dout, wt_packed, din, dtype, st);
}

int p2p_conv_wgrad(const p2p_conv* c, const void* in, const void* dout, float* dw_acc, int dtype, int impl,
                   void* stream) {
  if (int e = check_conv(c)) return e;
  P2P_REQUIRE(in && dout && dw_acc, "null pointer");
  P2P_REQUIRE(dtype == P2P_F32 || dtype == P2P_BF16, "unknown dtype");
  cudaStream_t st = as_stream(stream);
  const bool tc_ok = dtype == P2P_BF16 && device_tc() && conv_wgrad_tc_supported(c);
  P2P_REQUIRE(impl != P2P_IMPL_TC || tc_ok, "tcgen05 path does not support this shape / dtype / device");
  if (impl == P2P_IMPL_AUTO && conv_co1_supported(c)) return conv_wgrad_co1(c, in, dout, dw_acc, dtype, st);
  if (impl != P2P_IMPL_SIMT && tc_ok) return conv_wgrad_tc(c, in, dout, dw_acc, st);
  return conv_wgrad_simt(c, in, dout, dw_acc, dtype, st);
}

}  // extern "C"
