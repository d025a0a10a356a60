// abi_core.cu -- error string, version and device-query exports of libmgn_b200.so.
#include <stdarg.h>
#include <string.h>

#include "common.cuh"

static thread_local char g_err[512] = "";

void mgn_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

extern "C" const char* mgn_last_error(void) { return g_err; }

extern "C" int mgn_version(void) { return 100; /* 0.1.0 */ }

// 0 if the current device can run the sm_100a kernels, MGN_EUNSUPPORTED otherwise.
extern "C" int mgn_check_device(void) {
  int dev = 0;
  MGN_CUDA(cudaGetDevice(&dev));
  int major = 0, minor = 0;
  MGN_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
  MGN_CUDA(cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, dev));
  if (major != 10) {
    mgn_set_error("libmgn_b200 needs an sm_100a device (B200); found sm_%d%d", major, minor);
    return MGN_EUNSUPPORTED;
  }
  return MGN_OK;
}

// developer timeline buffer (see mlp_common.cuh Timeline); pass null to switch it off
long long* g_mgn_timeline_host = nullptr;
int g_mgn_timeline_cap_host = 0;
extern "C" int mgn_debug_set_timeline(long long* buf, int cap) {
  g_mgn_timeline_host = buf;
  g_mgn_timeline_cap_host = cap;
  return MGN_OK;
}
