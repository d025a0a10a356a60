// common.cuh -- error convention and small helpers shared by the C-ABI sources.
// Error convention (include/mgn_b200.h): every export returns int, 0 = ok,
// < 0 = argument / shape / unsupported, > 0 = a cudaError_t; never throws.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#define MGN_OK 0
#define MGN_EINVAL (-1)
#define MGN_EUNSUPPORTED (-2)
#define MGN_EWORKSPACE (-3)

void mgn_set_error(const char* fmt, ...);

#define MGN_CHECK_ARG(cond, ...)   \
  do {                             \
    if (!(cond)) {                 \
      mgn_set_error(__VA_ARGS__);  \
      return MGN_EINVAL;           \
    }                              \
  } while (0)

#define MGN_CUDA(call)                                                                   \
  do {                                                                                   \
    cudaError_t _e = (call);                                                             \
    if (_e != cudaSuccess) {                                                             \
      mgn_set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(_e), __FILE__, __LINE__); \
      return (int)_e;                                                                    \
    }                                                                                    \
  } while (0)

#define MGN_LAUNCH_CHECK()                                                               \
  do {                                                                                   \
    cudaError_t _e = cudaGetLastError();                                                 \
    if (_e != cudaSuccess) {                                                             \
      mgn_set_error("kernel launch failed: %s (%s:%d)", cudaGetErrorString(_e), __FILE__, __LINE__); \
      return (int)_e;                                                                    \
    }                                                                                    \
  } while (0)

static inline int mgn_num_sms() {
  static int sms = 0;
  if (sms == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (sms <= 0) sms = 148;
  }
  return sms;
}

static inline int64_t ceil_div64(int64_t a, int64_t b) { return (a + b - 1) / b; }

// developer timeline buffer (abi_core.cu); kernels receive it through their parameter struct
extern long long* g_mgn_timeline_host;
extern int g_mgn_timeline_cap_host;

// ---------------------------------------------------------------- programmatic dependent launch
// Every step kernel is launched with cudaLaunchAttributeProgrammaticStreamSerialization (MGN_PDL=0 switches it off):
// its CTAs may be scheduled as soon as every CTA of the previous kernel in the stream has passed pdl_launch() or has
// exited -- i.e. while the tail of that kernel still runs -- and they block in pdl_wait() until the previous kernel
// has completed and its writes are visible.  Everything before pdl_wait() (barrier init, TMEM allocation) therefore
// overlaps the previous kernel's last wave, and the launch latency between dependent kernels disappears.  A kernel
// must not touch global memory before pdl_wait().
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

#include <stdlib.h>
// ---------------------------------------------------------------- hang record (tc05.cuh: bounded mbarrier waits)
// 8 x u64 of host-mapped pinned memory, allocated once per process (abi_core.cu).  Each translation unit keeps the
// device pointer in its own __device__ variable, set on that unit's first launch.
void* mgn_hang_buffer();                       // host pointer == device pointer (UVA), or null if the allocation failed
bool mgn_no_alloc_scope();                     // abi_core.cu: the calling entry point must not allocate
void mgn_set_no_alloc_scope(bool on);
static __device__ unsigned long long* g_mgn_hang_dev = nullptr;

static inline cudaError_t mgn_register_hang_buffer(cudaStream_t stream) {
  static bool done = false;
  if (done) return cudaSuccess;
  void* h = mgn_hang_buffer();
  if (h == nullptr && mgn_no_alloc_scope()) return cudaSuccess;  // not created yet and may not be: register later
  cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(stream, &cs) != cudaSuccess) { (void)cudaGetLastError(); return cudaSuccess; }
  if (cs != cudaStreamCaptureStatusNone) return cudaSuccess;     // first launch under capture: try again later
  cudaError_t e = cudaMemcpyToSymbolAsync(g_mgn_hang_dev, &h, sizeof(h), 0, cudaMemcpyHostToDevice, stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(stream);       // `h` is a stack variable
  done = true;
  return e;
}

template <class P>
static inline cudaError_t mgn_launch(void (*kernel)(P), int grid, int block, size_t smem, cudaStream_t stream,
                                     const P& p) {
  static const bool pdl = !(getenv("MGN_PDL") && atoi(getenv("MGN_PDL")) == 0);
  cudaError_t he = mgn_register_hang_buffer(stream);
  if (he != cudaSuccess) return he;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)grid);
  cfg.blockDim = dim3((unsigned)block);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, p);
}
#endif
