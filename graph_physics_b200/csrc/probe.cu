// probe.cu -- developer micro-benchmarks of the SM-local resources the fused MLP kernels compete for (TMEM read / write
// throughput by number of active warps, product issue-to-completion latency).  Not on the product path; exported so
// that tools/probe.py can run them on the B200 and profiles/ can cite measured constants instead of guesses.
#define MGN_FILE_ID 4
#include "common.cuh"
#include "tc05.cuh"

using namespace tc05;

struct ProbeArgs {
  long long* out;     // [64] cycle counts
  int iters;
};

// 512 threads, 512 TMEM columns, two 32 KB operand tiles (garbage contents: only timing matters)
__global__ void __launch_bounds__(512, 1) probe_kernel(ProbeArgs p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  const int tid = threadIdx.x, warp = tid >> 5;
  for (int i = tid; i < 65536 / 4; i += 512) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;   // fp16 ones
  if (warp == 0) tmem_alloc<512>(&tmem_slot);
  if (tid == 0) {
    mbar_init(&bar, 1);
    mbar_init_fence();
  }
  fence_async_smem();
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tmem = tmem_slot;
  const uint32_t lane_off = (uint32_t)((warp & 3) * 32) << 16;
  const uint32_t c0 = (uint32_t)(warp >> 2) * 32;
  float v[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) v[j] = (float)(tid + j);
  float sink = 0.f;
  int slot = 0;
  // ---- tcgen05.ld.32x32b.x32 throughput with nw = 1, 4, 8, 16 active warps (each warp: its own lane quarter,
  //      column block (warp>>2)*32 + 128*(i&3))
  for (int nw = 1; nw <= 16; nw <<= 1) {
    if (nw == 2) continue;
    __syncthreads();
    const long long t0 = clock64();
    if (warp < nw) {
      for (int i = 0; i < p.iters; ++i) {
        tmem_ld32(tmem + lane_off + c0 + 128 * (i & 3), v);
        tmem_ld_wait();
        sink += v[0] + v[13] + v[31];
      }
    }
    __syncthreads();
    const long long t1 = clock64();
    if (tid == 0) p.out[slot] = t1 - t0;
    ++slot;
  }
  // ---- same without waiting after each load (4 loads in flight)
  for (int nw = 4; nw <= 16; nw <<= 2) {
    __syncthreads();
    const long long t0 = clock64();
    if (warp < nw) {
      for (int i = 0; i < p.iters; i += 2) {   // 4 x16 loads = 2 x32-equivalents per trip
        float a[16], b[16], c[16], d[16];
        tmem_ld16(tmem + lane_off + c0, a);
        tmem_ld16(tmem + lane_off + c0 + 16, b);
        tmem_ld16(tmem + lane_off + c0 + 128, c);
        tmem_ld16(tmem + lane_off + c0 + 144, d);
        tmem_ld_wait();
        sink += a[0] + b[7] + c[9] + d[15];
      }
    }
    __syncthreads();
    const long long t1 = clock64();
    if (tid == 0) p.out[slot] = t1 - t0;
    ++slot;
  }
  // ---- tcgen05.st.32x32b.x32 throughput, 16 warps
  {
    __syncthreads();
    const long long t0 = clock64();
    for (int i = 0; i < p.iters; ++i) {
      tmem_st32(tmem + lane_off + c0 + 128 * (i & 3), v);
    }
    tmem_st_wait();
    __syncthreads();
    const long long t1 = clock64();
    if (tid == 0) p.out[slot] = t1 - t0;
    ++slot;
  }
  // ---- product latency: 8 x (M128 N128 K16) + commit -> barrier observed by thread 0; then 2, 4 products back to back
  uint32_t phase = 0;
  for (int np = 1; np <= 4; np <<= 1) {
    __syncthreads();
    long long t0 = 0;
    if (tid == 0) {
      t0 = clock64();
      for (int q = 0; q < p.iters / 8 + 1; ++q) {
        for (int k = 0; k < np; ++k)
          umma_tile(tmem + 128 * (k & 3), smem_u32(smem), 128, 0, smem_u32(smem + 32768), 128, 0, 0, 128, 8, 0);
        mma_commit(&bar);
        mbar_wait(&bar, phase);
        phase ^= 1;
      }
      p.out[slot] = clock64() - t0;
    }
    ++slot;
  }
  // ---- M = 64 product (N128, K128)
  {
    __syncthreads();
    if (tid == 0) {
      const long long t0 = clock64();
      const uint32_t idesc = make_idesc(64, 128, 0, 0);
      for (int q = 0; q < p.iters / 8 + 1; ++q) {
        for (uint32_t ks = 0; ks < 8; ++ks)
          mma_f16(tmem, desc_kmajor(smem_u32(smem), 128, ks), desc_kmajor(smem_u32(smem + 32768), 128, ks), idesc, ks > 0);
        mma_commit(&bar);
        mbar_wait(&bar, phase);
        phase ^= 1;
      }
      p.out[slot] = clock64() - t0;
    }
    ++slot;
  }
  // ---- N = 256 product (M128, K128): 64 KB B operand = both tiles
  {
    __syncthreads();
    if (tid == 0) {
      const long long t0 = clock64();
      for (int q = 0; q < p.iters / 8 + 1; ++q) {
        umma_tile(tmem, smem_u32(smem), 128, 0, smem_u32(smem), 256, 0, 0, 256, 8, 0);
        mma_commit(&bar);
        mbar_wait(&bar, phase);
        phase ^= 1;
      }
      p.out[slot] = clock64() - t0;
    }
    ++slot;
  }
  // ---- products running while 16 warps read TMEM (do they slow each other down?)
  {
    __syncthreads();
    const long long t0 = clock64();
    if (tid == 0) {
      for (int q = 0; q < p.iters / 8 + 1; ++q) {
        umma_tile(tmem + 256, smem_u32(smem), 128, 0, smem_u32(smem + 32768), 128, 0, 0, 128, 8, 0);
        mma_commit(&bar);
        mbar_wait(&bar, phase);
        phase ^= 1;
      }
      p.out[slot] = clock64() - t0;
    } else if (warp >= 4) {       // 12 warps read columns 0..127 meanwhile
      for (int i = 0; i < p.iters; ++i) {
        tmem_ld32(tmem + lane_off + ((warp >> 2) - 1) * 32, v);
        tmem_ld_wait();
        sink += v[0] + v[13] + v[31];
      }
    }
    __syncthreads();
    const long long t1 = clock64();
    if (tid == 32 * 4) p.out[slot + 1] = t1 - t0;
    slot += 2;
  }
  if (sink == 1.2345f) p.out[63] = 1;
  fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc<512>(tmem);
}

// out_host[0..]: cycles for (iters loads) ld x32 nw=1,4,8,16 | no-wait nw=4,16 | st x32 nw=16 | (iters/8+1) x {1,2,4}
// products | M=64 product | N=256 product | product under 12 reading warps, total of that phase
extern "C" int mgn_probe_sm(long long* out_dev, int iters, void* stream) {
  MGN_CHECK_ARG(out_dev && iters > 0, "probe: bad args");
  ProbeArgs p{out_dev, iters};
  MGN_CUDA(cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
  probe_kernel<<<1, 512, 65536, (cudaStream_t)stream>>>(p);
  MGN_LAUNCH_CHECK();
  return MGN_OK;
}
