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
