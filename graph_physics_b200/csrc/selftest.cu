// selftest.cu -- one-tile tcgen05 product used by tests/test_gpu_umma.py to pin the
// shared-memory descriptor conventions of tc05.cuh (K-major / MN-major views of the
// same swizzled tile, the N=16 "ones" operand used for bias gradients, K=256 and
// N=256 shapes) against a CPU product before the fused kernels rely on them.
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
