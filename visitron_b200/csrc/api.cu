#include <stdlib.h>
// Library-wide state: error string, device checks.
#include <stdarg.h>
#include <string.h>

#include <atomic>

#include "common.cuh"

namespace vt {

static thread_local char g_err[512] = "";
static int g_sms = 0;
static std::atomic<int> g_reserved_sms{0};  // SMs left to an overlapped NCCL exchange by the persistent kernels (vt_reserve_sms)

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int cuda_fail(cudaError_t e, const char* what, const char* file, int line) {
  set_error("CUDA error %d (%s) at %s:%d: %s", (int)e, cudaGetErrorString(e), file, line, what);
  return VT_ERR_CUDA;
}

int num_sms() {
  if (g_sms == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&g_sms, cudaDevAttrMultiProcessorCount, dev);
    if (g_sms <= 0) g_sms = 148;
  }
  return g_sms;
}

// SMs a persistent kernel should occupy: while a gradient exchange is in flight its CTAs hold `reserved` SMs, and a
// persistent CTA queued behind one of them would start only when the exchange kernel retires (a whole-kernel tail).
bool pdl_enabled() {
  static const bool on = [] { const char* e = getenv("VT_PDL"); return e && e[0] == '1'; }();   // off unless VT_PDL=1, see common.cuh
  return on;
}
int usable_sms() {
  int sms = num_sms() - g_reserved_sms.load(std::memory_order_relaxed);
  return sms < 16 ? 16 : sms;
}

int resolve_driver_entry_points();
void count_launch(int n);

// Device-resident step state for sync-free / CUDA-graph steps (see the header): one process-wide seed source (forward runs on the
// caller's thread, backward on autograd's, both must see it) and a per-thread row limit (set and cleared around a group of launches).
static std::atomic<const uint32_t*> g_seed_source{nullptr};
static thread_local const int32_t* g_row_limit = nullptr;
const uint32_t* seed_source() { return g_seed_source.load(std::memory_order_relaxed); }
const int32_t* row_limit() { return g_row_limit; }
static thread_local uint32_t* g_keep_bits = nullptr;
uint32_t* attention_keep_bits() { return g_keep_bits; }
static thread_local const int32_t* g_bwd_sched = nullptr;
static thread_local int g_bwd_sched_ctas = 0;
const int32_t* attention_bwd_schedule(int* ctas) { *ctas = g_bwd_sched_ctas; return g_bwd_sched; }

// state[0] += golden ratio; state[1] = splitmix64(state[0]): one thread, first node of a captured step
__global__ void seed_advance_kernel(unsigned long long* state) {
  unsigned long long z = (state[0] += 0x9E3779B97F4A7C15ull);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  state[1] = z ^ (z >> 31);
}

}  // namespace vt

extern "C" const char* vt_last_error(void) { return vt::g_err; }

extern "C" int vt_num_sms(void) { return vt::num_sms(); }
extern "C" void vt_reserve_sms(int n) { vt::g_reserved_sms.store(n < 0 ? 0 : n); }

extern "C" void vt_set_attention_bwd_schedule(const int32_t* sched, int64_t ctas) { vt::g_bwd_sched = sched; vt::g_bwd_sched_ctas = sched ? (int)ctas : 0; }
extern "C" void vt_set_attention_keep_bits(uint32_t* bits) { vt::g_keep_bits = bits; }
extern "C" int64_t vt_attention_keep_bits_words(int64_t B, int64_t S, int64_t heads) { return B * heads * 4 * ((S + 127) / 128) * S; }
extern "C" void vt_set_seed_source(const uint64_t* seed_dev) { vt::g_seed_source.store(reinterpret_cast<const uint32_t*>(seed_dev)); }
extern "C" void vt_set_row_limit(const int32_t* rows_dev) { vt::g_row_limit = rows_dev; }
extern "C" int vt_seed_advance(uint64_t* state, vt_stream stream) {
  VT_REQUIRE(state && ((uintptr_t)state & 7) == 0, "vt_seed_advance: state must be an 8-byte aligned device uint64[2]");
  vt::seed_advance_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(reinterpret_cast<unsigned long long*>(state));
  vt::count_launch(1);
  VT_CHECK_LAUNCH();
  return VT_OK;
}

extern "C" int vt_init(int device) {
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0) {
    vt::set_error("no CUDA device (%s): libvisitron_b200 has no CPU fallback", cudaGetErrorString(e));
    return VT_ERR_NO_DEVICE;
  }
  VT_CUDA(cudaSetDevice(device));
  int major = 0, minor = 0;
  VT_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device));
  VT_CUDA(cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, device));
  if (major != 10) {
    vt::set_error("device %d is sm_%d%d; this library is built for sm_100a (B200) only", device, major, minor);
    return VT_ERR_NO_DEVICE;
  }
  vt::g_sms = 0;
  vt::num_sms();
  return vt::resolve_driver_entry_points();
}
