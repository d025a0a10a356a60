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
