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
  // ---- where does the issuing thread spend the product's time?  cycles in {8 x tcgen05.mma, tcgen05.commit, mbarrier wait}
  //      alone (out[14..16]) and while warps 4..15 stream shared memory (ld.shared.v4 of the 64 KB, out[17..19])
  for (int busy = 0; busy < 2; ++busy) {
    __syncthreads();
    long long ti = 0, tc = 0, tw = 0;
    if (tid == 0) {
      for (int q = 0; q < p.iters / 8 + 1; ++q) {
        const long long a0 = clock64();
        umma_tile(tmem + 256, smem_u32(smem), 128, 0, smem_u32(smem + 32768), 128, 0, 0, 128, 8, 0);
        const long long a1 = clock64();
        mma_commit(&bar);
        const long long a2 = clock64();
        mbar_wait(&bar, phase);
        phase ^= 1;
        const long long a3 = clock64();
        ti += a1 - a0; tc += a2 - a1; tw += a3 - a2;
      }
      p.out[slot] = ti; p.out[slot + 1] = tc; p.out[slot + 2] = tw;
    } else if (busy && warp >= 4) {
      for (int i = 0; i < p.iters * 4; ++i) {
        const uint4 u = *reinterpret_cast<const uint4*>(smem + ((tid * 16 + i * 8192) & 65535));
        sink += __uint_as_float(u.x) + __uint_as_float(u.w);
      }
    }
    slot += 3;
  }
  if (sink == 1.2345f) p.out[63] = 1;
  fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc<512>(tmem);
}

// out_host[0..]: cycles for (iters loads) ld x32 nw=1,4,8,16 | no-wait nw=4,16 | st x32 nw=16 | (iters/8+1) x {1,2,4}
// products | M=64 product | N=256 product | product under 12 reading warps, total of that phase
// ------------------------------------------------------------------------------------------------------------------
// SM <-> L2 / DRAM path, one 128-thread CTA per SM, 6 x 32 KB shared-memory slots, bulk copies issued by thread 0:
//   mode 0  g2s, every CTA reads the same 128 KB (L2 hits)          mode 1  g2s, every CTA streams its own region (DRAM)
//   mode 2  s2g, every CTA streams its own region                   mode 3  reduce-add f32, own region
//   mode 4  modes 1 + 2 together (3 slots each)
//   mode 5  mode 2, while warp 1 measures the latency of dependent 16-byte loads of a small L2-resident table
//   mode 6  the loads of mode 5 alone
//   mode 7  mode 4 on slots 2..5 while thread 64 runs M128 N128 K128 products on slots 0, 1 (does bulk-copy traffic
//           through shared memory slow the tensor core's operand fetch?); mode 8: the products alone
//   mode 9 / 10  the products, each preceded by a 64 KB reduce-add / plain bulk store issued by the same thread
//           out[block*4 + 2] = cycles in issue + commit, out[block*4 + 3] = cycles in the mbarrier wait (32 products)
// out[block*4 + {0,1,2}] = cycles of the CTA's loop, chunks moved, sum of the load latencies (modes 5, 6)
#include "mlp_common.cuh"
using namespace mgn;
struct MemProbeArgs {
  uint8_t* buf;        // >= grid * chunks * 32 KB (+ 128 KB)
  const uint4* table;  // 64 KB, L2-resident
  long long* out;
  int mode, chunks;
};
__global__ void __launch_bounds__(128, 1) mem_probe_kernel(MemProbeArgs p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bars[6];
  __shared__ uint64_t mbar_p;
  __shared__ uint32_t tmem_slot;
  __shared__ int stop;
  const int tid = threadIdx.x;
  constexpr uint32_t CH = 32768;
  if (tid == 0) {
    for (int i = 0; i < 6; ++i) mbar_init(&bars[i], 1);
    mbar_init(&mbar_p, 1);
    mbar_init_fence();
    stop = 0;
  }
  if (p.mode >= 7 && tid < 32) tmem_alloc<128>(&tmem_slot);
  fence_before_sync();
  for (int i = tid; i < 6 * (int)CH / 4; i += 128) reinterpret_cast<float*>(smem)[i] = 1.0f;
  fence_async_smem();
  __syncthreads();
  uint8_t* mine = p.buf + (size_t)blockIdx.x * p.chunks * CH;
  const uint64_t pol = policy_evict_first();
  const int mode = p.mode;
  fence_after_sync();
  const long long t0 = clock64();
  if (tid == 0 && mode != 6 && mode < 8) {
    const bool loads = mode == 0 || mode == 1 || mode == 4 || mode == 7;
    const bool stores = mode == 2 || mode == 3 || mode == 4 || mode == 5 || mode == 7;
    const int base = mode == 7 ? 2 : 0;                   // mode 7: slots 0, 1 are the products' operands
    const int nl = loads ? (stores ? (mode == 7 ? 2 : 3) : 6) : 0;   // slots [base, base+nl) take loads, the rest feed stores
    const int ns = stores ? 6 - base - nl : 0;
    uint32_t ph[6] = {0, 0, 0, 0, 0, 0};
    int li = 0, lw = 0, si = 0;
    while (lw < (loads ? p.chunks : 0) || si < (stores ? p.chunks : 0)) {
      if (loads) {
        while (li < p.chunks && li - lw < nl) {
          const int sl = li % nl;
          const uint8_t* src = mode == 0 ? p.buf + (size_t)(li & 3) * CH : mine + (size_t)li * CH;
          mbar_arrive_expect_tx(&bars[sl], CH);
          bulk_g2s(smem + (base + sl) * CH, src, CH, &bars[sl]);
          ++li;
        }
        if (lw < li) {
          const int sl = lw % nl;
          mbar_wait(&bars[sl], ph[sl]);
          ph[sl] ^= 1;
          ++lw;
        }
      }
      if (stores && si < p.chunks) {
        uint8_t* dst = mine + (size_t)(mode == 4 ? p.chunks - 1 - si : si) * CH;
        if (mode == 3) bulk_red_add_f32(reinterpret_cast<float*>(dst), smem + (base + nl + si % ns) * CH, CH, pol);
        else bulk_s2g(dst, smem + (base + nl + si % ns) * CH, CH, pol);
        bulk_commit();
        ++si;
        if (ns == 2) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
        else if (ns == 3) asm volatile("cp.async.bulk.wait_group.read 2;" ::: "memory");
        else asm volatile("cp.async.bulk.wait_group.read 5;" ::: "memory");
      }
    }
    bulk_wait0();
    stop = 1;
  }
  long long lat = 0;
  int nlat = 0;
  if (tid == 64 && mode >= 7) {
    uint32_t ph = 0;
    long long ti = 0, tw = 0;
    for (int q = 0; q < 32; ++q) {
      if (mode == 9 || mode == 10) {          // 9: a 64 KB reduce-add / 10: a 64 KB plain store, issued by THIS thread first
        uint8_t* dst = mine + (size_t)(q & 63) * 2 * CH;
        if (mode == 9) bulk_red_add_f32(reinterpret_cast<float*>(dst), smem + 2 * CH, 2 * CH, pol);
        else bulk_s2g(dst, smem + 2 * CH, 2 * CH, pol);
        bulk_commit();
      }
      const long long a0 = clock64();
      umma_tile(tmem_slot, smem_u32(smem), 128, 0, smem_u32(smem + CH), 128, 0, 0, 128, 8, 0);
      mma_commit(&mbar_p);
      const long long a1 = clock64();
      mbar_wait(&mbar_p, ph);
      ph ^= 1;
      const long long a2 = clock64();
      ti += a1 - a0; tw += a2 - a1;
      if (mode == 9 || mode == 10) bulk_wait_read0();
    }
    if (mode == 9 || mode == 10) bulk_wait0();
    p.out[blockIdx.x * 4 + 2] = ti;
    p.out[blockIdx.x * 4 + 3] = tw;
  }
  if (tid == 32 && (mode == 5 || mode == 6)) {
    uint32_t idx = blockIdx.x * 97u;
    volatile int* vstop = &stop;
    while (mode == 6 ? nlat < 4096 : !*vstop) {
      const long long a = clock64();
      const uint4 v = __ldcg(p.table + (idx & 4095u));
      idx = idx * 1664525u + 1013904223u + (v.x & 1u);
      const long long b = clock64();
      lat += b - a;
      ++nlat;
    }
  }
  __syncthreads();
  const long long t1 = clock64();
  if (tid == 0) {
    p.out[blockIdx.x * 4 + 0] = t1 - t0;
    p.out[blockIdx.x * 4 + 1] = p.chunks;
  }
  if (tid == 32 && mode < 7) {
    p.out[blockIdx.x * 4 + 2] = lat;
    p.out[blockIdx.x * 4 + 3] = nlat;
  }
  if (mode >= 7) {
    fence_before_sync();
    __syncthreads();
    if (tid < 32) tmem_dealloc<128>(tmem_slot);
  }
}

extern "C" int mgn_probe_mem(void* buf, const void* table, long long* out_dev, int mode, int chunks, int grid,
                             void* stream) {
  MGN_CHECK_ARG(buf && table && out_dev && chunks > 0 && grid > 0 && mode >= 0 && mode <= 10, "probe_mem: bad args");
  MemProbeArgs p{(uint8_t*)buf, (const uint4*)table, out_dev, mode, chunks};
  MGN_CUDA(cudaFuncSetAttribute(mem_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 6 * 32768));
  mem_probe_kernel<<<grid, 128, 6 * 32768, (cudaStream_t)stream>>>(p);
  MGN_LAUNCH_CHECK();
  return MGN_OK;
}

extern "C" int mgn_probe_sm(long long* out_dev, int iters, void* stream) {
  MGN_CHECK_ARG(out_dev && iters > 0, "probe: bad args");
  ProbeArgs p{out_dev, iters};
  MGN_CUDA(cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
  probe_kernel<<<1, 512, 65536, (cudaStream_t)stream>>>(p);
  MGN_LAUNCH_CHECK();
  return MGN_OK;
}
