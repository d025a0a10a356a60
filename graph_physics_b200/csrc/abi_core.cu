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

extern "C" int mgn_version(void) { return 200; /* 0.2.0: round-2 ABI (mgn_edge_fwd removed; carry / rowptr / norm_dim arguments; chained entry points) */ }

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

// ---------------------------------------------------------------- hang record (tc05.cuh: bounded mbarrier waits)
// Host-mapped pinned memory, so that the record survives the trap that follows it (device memory of a faulted context
// cannot be read back).  Layout (u64): [0] magic, [1] tag = file id << 16 | line, [2] block << 32 | thread,
// [3] barrier shared address << 32 | parity, [4] grid size << 32 | SM id.
static unsigned long long* g_hang_host = nullptr;
static bool g_hang_tried = false;

// set (per host thread) by entry points that promise not to allocate: the hang buffer is then used if it already
// exists, but not created
thread_local bool g_mgn_no_alloc_scope = false;
bool mgn_no_alloc_scope() { return g_mgn_no_alloc_scope; }
void mgn_set_no_alloc_scope(bool on) { g_mgn_no_alloc_scope = on; }

void* mgn_hang_buffer() {
  if (!g_hang_tried && g_mgn_no_alloc_scope) return nullptr;      // would have to cudaHostAlloc: not in this scope
  if (!g_hang_tried) {
    g_hang_tried = true;
    void* h = nullptr;
    if (cudaHostAlloc(&h, 8 * sizeof(unsigned long long), cudaHostAllocMapped | cudaHostAllocPortable) == cudaSuccess) {
      memset(h, 0, 8 * sizeof(unsigned long long));
      g_hang_host = (unsigned long long*)h;
    } else {
      (void)cudaGetLastError();
    }
  }
  return g_hang_host;
}

static const char* hang_file(unsigned id) {
  switch (id) {
    case 1: return "mlp_fwd.cu";
    case 2: return "edge_fwd_hl.cu";
    case 3: return "mlp_bwd.cu";
    case 4: return "selftest.cu";
    default: return "?";
  }
}

// Writes a one-line description of the recorded mbarrier time-out into buf (empty string if none); returns 1 if a
// record exists.  Safe to call after a failed launch / synchronize: it only reads host memory.
extern "C" int mgn_hang_report(char* buf, int n) {
  if (buf && n > 0) buf[0] = 0;
  const unsigned long long* h = g_hang_host;
  if (h == nullptr || h[0] == 0) return 0;
  if (buf && n > 0)
    snprintf(buf, (size_t)n,
             "mbarrier wait timed out: %s:%u block %u of %u thread %u sm %u barrier@smem 0x%x parity %u",
             hang_file((unsigned)(h[1] >> 16)), (unsigned)(h[1] & 0xffff), (unsigned)(h[2] >> 32), (unsigned)(h[4] >> 32),
             (unsigned)(h[2] & 0xffffffffu), (unsigned)(h[4] & 0xffffffffu), (unsigned)(h[3] >> 32),
             (unsigned)(h[3] & 0xffffffffu));
  return 1;
}
