// selftest.cu -- one-tile tcgen05 product used by tests/test_gpu_umma.py to pin the
// shared-memory descriptor conventions of tc05.cuh (K-major / MN-major views of the
// same swizzled tile, the N=16 "ones" operand used for bias gradients, K=256 and
// N=256 shapes) against a CPU product before the fused kernels rely on them.
#define MGN_FILE_ID 4
#include "common.cuh"
#include "tc05.cuh"

using namespace tc05;

struct SelftestArgs {
  const __half* a;   // [128][a_cols] row-major
  const __half* b;   // [128][b_cols] row-major
  float* d;          // [128][n]
  int a_cols, b_cols;
  int a_mn, b_mn;    // 0 = K-major view, 1 = MN-major view
  int n;             // UMMA N
  int ksteps;        // 16-wide k steps
  int b_col0;        // MN-major B: first column
  int use_ones;      // B = 16-column all-ones operand
  int repeat;        // issue the whole product `repeat` times, accumulating
  int lbo_a, sbo_a, lbo_b, sbo_b;  // >=0: override descriptor strides (bytes)
};

__global__ void __launch_bounds__(128, 1) umma_selftest_kernel(SelftestArgs p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* tileA = smem;                  // up to 64 KB
  uint8_t* tileB = smem + 65536;          // up to 64 KB
  uint8_t* ones = smem + 131072;          // 512 B
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  const int tid = threadIdx.x, warp = tid >> 5;

  for (int i = tid; i < 128 * p.a_cols; i += 128) {
    int r = i / p.a_cols, c = i % p.a_cols;
    *reinterpret_cast<__half*>(tileA + tile_off(128, r, c)) = p.a[i];
  }
  for (int i = tid; i < 128 * p.b_cols; i += 128) {
    int r = i / p.b_cols, c = i % p.b_cols;
    *reinterpret_cast<__half*>(tileB + tile_off(128, r, c)) = p.b[i];
  }
  for (int i = tid; i < 256; i += 128) reinterpret_cast<__half*>(ones)[i] = __float2half(1.0f);
  if (warp == 0) tmem_alloc<256>(&tmem_slot);
  if (tid == 0) {
    mbar_init(&bar, 1);
    mbar_init_fence();
  }
  fence_async_smem();
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tmem = tmem_slot;

  if (tid == 0) {
    const uint32_t idesc = make_idesc(128, p.n, p.a_mn, p.b_mn);
    for (int rep = 0; rep < p.repeat; ++rep) {
      for (int ks = 0; ks < p.ksteps; ++ks) {
        uint64_t ad = p.a_mn ? desc_mnmajor(smem_u32(tileA), 128, ks) : desc_kmajor(smem_u32(tileA), 128, ks);
        uint64_t bd;
        if (p.use_ones) {
          bd = make_desc(smem_u32(ones), 128, 256, LAYOUT_NONE);
        } else {
          bd = p.b_mn ? desc_mnmajor(smem_u32(tileB), 128, ks, p.b_col0) : desc_kmajor(smem_u32(tileB), 128, ks);
        }
        if (p.lbo_a >= 0) ad = (ad & ~(0x3FFFull << 16)) | ((uint64_t)((p.lbo_a >> 4) & 0x3FFF) << 16);
        if (p.sbo_a >= 0) ad = (ad & ~(0x3FFFull << 32)) | ((uint64_t)((p.sbo_a >> 4) & 0x3FFF) << 32);
        if (p.lbo_b >= 0) bd = (bd & ~(0x3FFFull << 16)) | ((uint64_t)((p.lbo_b >> 4) & 0x3FFF) << 16);
        if (p.sbo_b >= 0) bd = (bd & ~(0x3FFFull << 32)) | ((uint64_t)((p.sbo_b >> 4) & 0x3FFF) << 32);
        mma_f16(tmem, ad, bd, idesc, (rep > 0 || ks > 0) ? 1u : 0u);
      }
    }
    mma_commit(&bar);
  }
  mbar_wait(&bar, 0);
  fence_after_sync();

  const uint32_t lane_base = tmem + ((uint32_t)(warp * 32) << 16);
  for (int c0 = 0; c0 < p.n; c0 += 16) {
    float v[16];
    tmem_ld16(lane_base + c0, v);
    tmem_ld_wait();
    for (int j = 0; j < 16; ++j) p.d[(size_t)tid * p.n + c0 + j] = v[j];
  }
  fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc<256>(tmem);
}

extern "C" int mgn_selftest_umma(const void* a, const void* b, float* d, int a_cols, int b_cols, int a_mn,
                                 int b_mn, int n, int ksteps, int b_col0, int use_ones, int repeat,
                                 int lbo_a, int sbo_a, int lbo_b, int sbo_b, void* stream) {
  MGN_CHECK_ARG(a && b && d, "selftest: null pointer");
  MGN_CHECK_ARG((a_cols == 128 || a_cols == 256) && (b_cols == 128 || b_cols == 256), "selftest: cols");
  MGN_CHECK_ARG(n >= 16 && n <= 256 && n % 16 == 0, "selftest: n");
  SelftestArgs p{(const __half*)a, (const __half*)b, d, a_cols, b_cols, a_mn, b_mn, n, ksteps, b_col0,
                 use_ones, repeat, lbo_a, sbo_a, lbo_b, sbo_b};
  const int smem = 131072 + 512 + 1024;
  MGN_CUDA(cudaFuncSetAttribute(umma_selftest_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  umma_selftest_kernel<<<1, 128, smem, (cudaStream_t)stream>>>(p);
  MGN_LAUNCH_CHECK();
  return MGN_OK;
}

// ---------------------------------------------------------------------------------
// Probe 2 (round-2 groundwork): variants the fused kernels do not use yet.
//   m = 64            : UMMA M = 64 (which TMEM lanes receive the 64 accumulator rows)
//   a_tmem = 1 / 2    : A operand read from TENSOR MEMORY (tcgen05.mma [d], [a], b-desc): the A tile is written with
//                       tcgen05.st, 16 K-elements = 8 32-bit columns per k-step; 1 = element 2c in the low half of
//                       column c, 2 = in the high half
// d receives all 128 TMEM lanes x n columns, whatever the shape, so the host can see where rows land.
struct Selftest2Args {
  const __half* a;   // [128][128] row-major, K-major operand
  const __half* b;   // [128][128] row-major, K-major operand
  float* d;          // [128][n]
  int n, ksteps, m, a_tmem;
};

__device__ __forceinline__ void tmem_st8(uint32_t taddr, const uint32_t* r) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"r"(taddr), "r"(r[0]),
               "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7])
               : "memory");
}

__global__ void __launch_bounds__(128, 1) umma_selftest2_kernel(Selftest2Args p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* tileA = smem;
  uint8_t* tileB = smem + 32768;
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  const int tid = threadIdx.x, warp = tid >> 5;
  for (int i = tid; i < 128 * 128; i += 128) {
    int r = i >> 7, c = i & 127;
    *reinterpret_cast<__half*>(tileA + tile_off(128, r, c)) = p.a[i];
    *reinterpret_cast<__half*>(tileB + tile_off(128, r, c)) = p.b[i];
  }
  if (warp == 0) tmem_alloc<256>(&tmem_slot);
  if (tid == 0) {
    mbar_init(&bar, 1);
    mbar_init_fence();
  }
  fence_async_smem();
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tmem = tmem_slot;
  const uint32_t lane_base = tmem + ((uint32_t)(warp * 32) << 16);
  const uint32_t a_col0 = 128;            // A operand columns [128, 192)
  // clear the accumulator lanes so that untouched lanes read as zero
  {
    uint32_t z[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int c = 0; c < 128; c += 8) tmem_st8(lane_base + c, z);
    tmem_st_wait();
  }
  if (p.a_tmem) {   // thread t owns TMEM lane t = row t of A
    for (int ks = 0; ks < 8; ++ks) {
      uint32_t r[8];
      for (int c = 0; c < 8; ++c) {
        const uint32_t lo = __half_as_ushort(p.a[tid * 128 + ks * 16 + 2 * c]);
        const uint32_t hi = __half_as_ushort(p.a[tid * 128 + ks * 16 + 2 * c + 1]);
        r[c] = p.a_tmem == 1 ? (lo | (hi << 16)) : (hi | (lo << 16));
      }
      tmem_st8(lane_base + a_col0 + ks * 8, r);
    }
    tmem_st_wait();
  }
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  if (tid == 0) {
    const uint32_t idesc = make_idesc(p.m, p.n, 0, 0);
    for (int ks = 0; ks < p.ksteps; ++ks) {
      const uint64_t bd = desc_kmajor(smem_u32(tileB), 128, ks);
      if (p.a_tmem) mma_f16_ts(tmem, tmem + a_col0 + ks * 8, bd, idesc, ks > 0 ? 1u : 0u);
      else mma_f16(tmem, desc_kmajor(smem_u32(tileA), 128, ks), bd, idesc, ks > 0 ? 1u : 0u);
    }
    mma_commit(&bar);
  }
  mbar_wait(&bar, 0);
  fence_after_sync();
  for (int c0 = 0; c0 < p.n; c0 += 16) {
    float v[16];
    tmem_ld16(lane_base + c0, v);
    tmem_ld_wait();
    for (int j = 0; j < 16; ++j) p.d[(size_t)tid * p.n + c0 + j] = v[j];
  }
  fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc<256>(tmem);
}

extern "C" int mgn_selftest_umma2(const void* a, const void* b, float* d, int n, int ksteps, int m, int a_tmem,
                                  void* stream) {
  MGN_CHECK_ARG(a && b && d, "selftest2: null pointer");
  MGN_CHECK_ARG(n >= 16 && n <= 128 && n % 16 == 0 && ksteps >= 1 && ksteps <= 8, "selftest2: n / ksteps");
  MGN_CHECK_ARG((m == 64 || m == 128) && a_tmem >= 0 && a_tmem <= 2, "selftest2: m / a_tmem");
  Selftest2Args p{(const __half*)a, (const __half*)b, d, n, ksteps, m, a_tmem};
  const int smem = 65536 + 1024;
  MGN_CUDA(cudaFuncSetAttribute(umma_selftest2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  umma_selftest2_kernel<<<1, 128, smem, (cudaStream_t)stream>>>(p);
  MGN_LAUNCH_CHECK();
  return MGN_OK;
}

// ------------------------------------------------------------------------------------------------ CTA pair probe
// One tcgen05.mma.cta_group::2 product D[256 x N] = A[256 x 16 ksteps] * B[N x 16 ksteps]^T on a cluster of two CTAs
// (the shape DESIGN.md plans for the round-2 edge kernels): CTA r holds rows 128 r .. 128 r + 127 of A (shared memory
// tile, or tensor memory when a_tmem) and rows N/2 * r .. of B as an (N/2)-row panel tile -- half of every weight tile
// per CTA; the leader's elected thread issues the instruction stream for both, one commit is multicast to both CTAs'
// barriers, each CTA reads its own 128 accumulator lanes.  Probe only: NOT used by the fused kernels, and not run by
// the default test suite until it has been confirmed on a B200 (tests/test_gpu_umma.py, MGN_TEST_PAIR=1).
struct SelftestPairArgs {
  const __half* a;   // [256][128]
  const __half* b;   // [n][128]
  float* d;          // [256][n]
  int n, ksteps, a_tmem;
};

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128, 1) umma_selftest_pair_kernel(SelftestPairArgs p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* tileA = smem;                  // 128 rows x 128 fp16
  uint8_t* tileB = smem + 32768;          // n/2 rows x 128 fp16 (panel pitch = n/2 rows)
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  const int tid = threadIdx.x, warp = tid >> 5;
  const uint32_t rank = cluster_ctarank();
  const int nh = p.n >> 1;
  for (int i = tid; i < 128 * 128; i += 128) {
    const int r = i >> 7, c = i & 127;
    *reinterpret_cast<__half*>(tileA + tile_off(128, r, c)) = p.a[(size_t)(128 * rank + r) * 128 + c];
  }
  for (int i = tid; i < nh * 128; i += 128) {
    const int r = i >> 7, c = i & 127;
    *reinterpret_cast<__half*>(tileB + tile_off(nh, r, c)) = p.b[(size_t)(nh * rank + r) * 128 + c];
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_slot)),
                 "n"(256)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  if (tid == 0) {
    mbar_init(&bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  fence_async_smem();
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tmem = tmem_slot;
  const uint32_t lane_base = tmem + ((uint32_t)(warp * 32) << 16);
  const uint32_t a_col0 = 128;
  if (p.a_tmem) {   // thread t owns TMEM lane t = row 128 rank + t of A
    for (int ks = 0; ks < 8; ++ks) {
      uint32_t r[8];
      for (int c = 0; c < 8; ++c) {
        const size_t o = (size_t)(128 * rank + tid) * 128 + ks * 16 + 2 * c;
        r[c] = (uint32_t)__half_as_ushort(p.a[o]) | ((uint32_t)__half_as_ushort(p.a[o + 1]) << 16);
      }
      tmem_st8(lane_base + a_col0 + ks * 8, r);
    }
    tmem_st_wait();
  }
  fence_before_sync();
  cluster_sync_all();          // both CTAs' operands (shared and tensor memory) and barriers are ready
  fence_after_sync();
  if (rank == 0 && tid == 0) {
    const uint32_t idesc = make_idesc(256, p.n, 0, 0);
    for (int ks = 0; ks < p.ksteps; ++ks) {
      const uint64_t bd = desc_kmajor(smem_u32(tileB), nh, ks);
      const uint32_t acc = ks > 0 ? 1u : 0u;
      if (p.a_tmem) {
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                     "tcgen05.mma.cta_group::2.kind::f16 [%0], [%1], %2, %3, p;\n\t}\n" ::"r"(tmem),
                     "r"(tmem + a_col0 + ks * 8), "l"(bd), "r"(idesc), "r"(acc)
                     : "memory");
      } else {
        const uint64_t ad = desc_kmajor(smem_u32(tileA), 128, ks);
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                     "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(tmem),
                     "l"(ad), "l"(bd), "r"(idesc), "r"(acc)
                     : "memory");
      }
    }
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::
                     "r"(smem_u32(&bar)), "h"((uint16_t)0b11)
                 : "memory");
  }
  mbar_wait(&bar, 0);
  fence_after_sync();
  for (int c0 = 0; c0 < p.n; c0 += 16) {
    float v[16];
    tmem_ld16(lane_base + c0, v);
    tmem_ld_wait();
    for (int j = 0; j < 16; ++j) p.d[(size_t)(128 * rank + tid) * p.n + c0 + j] = v[j];
  }
  fence_before_sync();
  cluster_sync_all();          // neither CTA frees tensor memory (or exits) while the other still reads
  if (warp == 0)
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(256) : "memory");
}

extern "C" int mgn_selftest_umma_pair(const void* a, const void* b, float* d, int n, int ksteps, int a_tmem,
                                      void* stream) {
  MGN_CHECK_ARG(a && b && d, "selftest_pair: null pointer");
  MGN_CHECK_ARG(n >= 32 && n <= 256 && n % 32 == 0 && ksteps >= 1 && ksteps <= 8, "selftest_pair: n / ksteps");
  MGN_CHECK_ARG(a_tmem == 0 || a_tmem == 1, "selftest_pair: a_tmem");
  SelftestPairArgs p{(const __half*)a, (const __half*)b, d, n, ksteps, a_tmem};
  const int smem = 65536 + 1024;
  MGN_CUDA(cudaFuncSetAttribute(umma_selftest_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  umma_selftest_pair_kernel<<<2, 128, smem, (cudaStream_t)stream>>>(p);
  MGN_LAUNCH_CHECK();
  return MGN_OK;
}
