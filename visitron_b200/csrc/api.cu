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

}  // namespace vt

extern "C" const char* vt_last_error(void) { return vt::g_err; }

extern "C" int vt_num_sms(void) { return vt::num_sms(); }
extern "C" void vt_reserve_sms(int n) { vt::g_reserved_sms.store(n < 0 ? 0 : n); }

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
