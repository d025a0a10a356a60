// tc05.cuh -- thin inline-PTX layer over the sm_100a tensor-core path used by the
// MeshGraphNet kernels: tcgen05.mma (kind::f16, fp16 operands, fp32 accumulate in
// TMEM), TMEM alloc / ld, mbarrier, proxy fences, and the shared-memory tile
// layout every kernel in this directory uses.
//
// Tile layout ("panel tile"): an fp16 matrix of R rows x C columns (C multiple of
// 64) is stored as C/64 panels; one panel = R rows x 64 columns, row pitch 128 B,
// 128-byte swizzle (the 16-byte chunk index of a row is XORed with row % 8).
// Panels are R*128 bytes apart.  The base must be 1024-byte aligned.
//   * read as a K-major UMMA operand:  rows are the M (or N) index, columns are K
//   * read as an MN-major UMMA operand: rows are the K index, columns are M (or N)
// The bytes are identical in both views (canonical SWIZZLE_128B layouts of
// cute::UMMA, K-major ((8,n),2):((8,SBO),1) and MN-major ((8,n),(8,k)):((1,LBO),(8,SBO))
// in 16-byte units), so an activation tile written once serves as the A operand
// of the next layer (K-major) and as an operand of the weight-gradient product
// (MN-major) without being moved.
#pragma once
#include <cuda_fp16.h>
#include <stdint.h>

#include "common.cuh"

namespace tc05 {

static constexpr uint32_t PANEL_COLS = 64;
static constexpr uint32_t ROW_BYTES = 128;

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// byte offset of element (row, col) inside a panel tile with `rows` rows
__device__ __forceinline__ uint32_t tile_off(uint32_t rows, uint32_t row, uint32_t col) {
  uint32_t panel = col >> 6, cc = col & 63;
  uint32_t chunk = (cc >> 3) ^ (row & 7);
  return panel * rows * ROW_BYTES + row * ROW_BYTES + (chunk << 4) + ((cc & 7) << 1);
}

// ---------------------------------------------------------------- descriptors
// 64-bit shared-memory matrix descriptor (cute::UMMA::SmemDescriptor):
//  [0,14) start>>4  [16,30) LBO>>4  [32,46) SBO>>4  [46,48) version=1  [61,64) layout
static constexpr uint64_t LAYOUT_NONE = 0, LAYOUT_SW128 = 2;

__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes,
                                              uint64_t layout) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  d |= layout << 61;
  return d;
}

// K-major operand, k-step `ks` (16 columns per step) of a panel tile with `rows` rows.
__device__ __forceinline__ uint64_t desc_kmajor(uint32_t tile_saddr, uint32_t rows, uint32_t ks) {
  uint32_t a = tile_saddr + (ks >> 2) * rows * ROW_BYTES + (ks & 3) * 32;
  return make_desc(a, 16, 1024, LAYOUT_SW128);
}
// MN-major operand, k-step `ks` (16 rows per step); `rows` = total rows of the tile
// (panel pitch), `col0` = first MN column (multiple of 64).
__device__ __forceinline__ uint64_t desc_mnmajor(uint32_t tile_saddr, uint32_t rows, uint32_t ks,
                                                 uint32_t col0 = 0) {
  uint32_t a = tile_saddr + (col0 >> 6) * rows * ROW_BYTES + ks * 16 * ROW_BYTES;
  return make_desc(a, rows * ROW_BYTES, 1024, LAYOUT_SW128);
}

// 32-bit instruction descriptor (cute::UMMA::InstrDescriptor), kind::f16:
//  [4,6) D fmt (1=f32) [7,10) A fmt (0=f16) [10,13) B fmt [15] A major [16] B major
//  [17,23) N>>3 [24,29) M>>4 ; major: 0 = K, 1 = MN
__host__ __device__ constexpr uint32_t make_idesc(uint32_t M, uint32_t N, uint32_t a_mn, uint32_t b_mn) {
  return (1u << 4) | (a_mn << 15) | (b_mn << 16) | ((N >> 3) << 17) | ((M >> 4) << 24);
}

// ---------------------------------------------------------------- instructions
__device__ __forceinline__ void mma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                        uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive on an mbarrier when all previously issued MMAs of this thread are done
__device__ __forceinline__ void mma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                   smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void fence_before_sync() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void fence_after_sync() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
// generic-proxy smem writes -> visible to the async proxy (UMMA operand reads)
__device__ __forceinline__ void fence_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_init_fence() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
// ---- bounded mbarrier wait.  A lost arrive must become a CUDA error, not an endless spin: after MBAR_FAST_POLLS
// failed polls the waiter reads %globaltimer on every further poll and, once MBAR_TIMEOUT_NS have passed, writes a record
// {magic, tag (source file id << 16 | line of the wait), block, thread, barrier address, parity, SM} to the host-mapped
// hang buffer (common.cuh: g_mgn_hang_dev, registered by mgn_launch) and traps.  The launch then fails with
// cudaErrorLaunchFailure and mgn_hang_report() / mgn_last_error() name the wait that starved.
static constexpr uint32_t MBAR_FAST_POLLS = 4096;
static constexpr unsigned long long MBAR_TIMEOUT_NS = 4000000000ull;   // 4 s: a legitimate wait is microseconds
static constexpr unsigned long long HANG_MAGIC = 0x4d474e48414e4721ull;  // "MGNHANG!"

static __device__ __noinline__ void mbar_timeout(uint32_t bar_saddr, uint32_t parity, uint32_t tag) {
  unsigned long long* h = g_mgn_hang_dev;
  if (h != nullptr && atomicCAS(h, 0ull, HANG_MAGIC) == 0ull) {      // first reporter wins
    uint32_t smid;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    h[1] = tag;
    h[2] = ((unsigned long long)blockIdx.x << 32) | threadIdx.x;
    h[3] = ((unsigned long long)bar_saddr << 32) | parity;
    h[4] = ((unsigned long long)gridDim.x << 32) | smid;
    __threadfence_system();
  }
  __trap();
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t a, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}\n"
      : "=r"(ok) : "r"(a), "r"(parity) : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait_tagged(uint64_t* bar, uint32_t parity, uint32_t tag) {
  const uint32_t a = smem_u32(bar);
#pragma unroll 1
  for (uint32_t i = 0; i < MBAR_FAST_POLLS; ++i)
    if (mbar_try_wait(a, parity)) return;
  unsigned long long t0;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
#pragma unroll 1
  while (!mbar_try_wait(a, parity)) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    if (t - t0 > MBAR_TIMEOUT_NS) mbar_timeout(a, parity, tag);
  }
}
#ifndef MGN_FILE_ID
#define MGN_FILE_ID 0
#endif
// every wait carries its source position: (file id << 16) | line
#define mbar_wait(bar, parity) mbar_wait_tagged((bar), (parity), ((uint32_t)(MGN_FILE_ID) << 16) | (uint32_t)__LINE__)
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
// 1-D bulk copy global -> shared, completion on an mbarrier (bytes multiple of 16)
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
          smem_u32(smem_dst)),
      "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}

// TMEM allocation: one full warp executes; result (base address) lands in *smem_slot.
template <uint32_t NCOLS>
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_slot) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_slot)),
               "n"(NCOLS)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
template <uint32_t NCOLS>
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(NCOLS) : "memory");
}

// TMEM -> registers: the calling warp reads its own 32 lanes (lane = 32*(warp%4)+laneid),
// 32 consecutive fp32 columns starting at `taddr` (lane field must already be set).
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float* v) {
  uint32_t* r = reinterpret_cast<uint32_t*>(v);
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
      "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]),
        "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]),
        "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]),
        "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float* v) {
  uint32_t* r = reinterpret_cast<uint32_t*>(v);
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]),
        "=r"(r[15])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() {
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// registers -> TMEM: the calling warp writes 16 consecutive 32-bit columns of its own 32 lanes
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t* r) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
      "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
      "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const float* v) {
  const uint32_t* r = reinterpret_cast<const uint32_t*>(v);
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,"
      "%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
      "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]),
      "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]),
      "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
      : "memory");
}
// one / four 32-bit columns of the calling warp's own 32 lanes
__device__ __forceinline__ void tmem_st1(uint32_t taddr, uint32_t r) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x1.b32 [%0], {%1};" ::"r"(taddr), "r"(r) : "memory");
}
__device__ __forceinline__ void tmem_ld4(uint32_t taddr, float* v) {
  uint32_t* r = reinterpret_cast<uint32_t*>(v);
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// A operand from tensor memory (pinned by tests/test_gpu_umma.py::test_a_operand_from_tensor_memory): row r of A is
// TMEM lane r; one k-step (16 fp16) is 8 consecutive 32-bit columns; K element 2c sits in the LOW half of column c.
__device__ __forceinline__ void mma_f16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc,
                                           uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}\n" ::"r"(tmem_d),
      "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}

// sub-block barrier over `nthreads` threads (ids 1..15; 0 is __syncthreads)
__device__ __forceinline__ void named_bar(uint32_t id, uint32_t nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

__device__ __forceinline__ uint32_t pack_h2(float a, float b) {
  __half2 h = __floats2half2_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}

// Issue one [128 x N] += A[128 x 16*ksteps] * B product, operands in panel tiles.
//   a_mn / b_mn: 0 = K-major view, 1 = MN-major view of the tile.
// Called by ONE thread.
__device__ __forceinline__ void umma_tile(uint32_t tmem_d, uint32_t a_saddr, uint32_t a_rows, uint32_t a_mn,
                                          uint32_t b_saddr, uint32_t b_rows, uint32_t b_mn, uint32_t b_col0,
                                          uint32_t N, uint32_t ksteps, uint32_t accumulate_first) {
  const uint32_t idesc = make_idesc(128, N, a_mn, b_mn);
  for (uint32_t ks = 0; ks < ksteps; ++ks) {
    uint64_t ad = a_mn ? desc_mnmajor(a_saddr, a_rows, ks) : desc_kmajor(a_saddr, a_rows, ks);
    uint64_t bd = b_mn ? desc_mnmajor(b_saddr, b_rows, ks, b_col0) : desc_kmajor(b_saddr, b_rows, ks);
    mma_f16(tmem_d, ad, bd, idesc, (ks > 0) ? 1u : accumulate_first);
  }
}

// Same product with the A operand (128 rows x 16*ksteps fp16) in tensor memory at column `tmem_a`.
__device__ __forceinline__ void umma_tile_ts(uint32_t tmem_d, uint32_t tmem_a, uint32_t b_saddr, uint32_t b_rows,
                                             uint32_t b_mn, uint32_t b_col0, uint32_t N, uint32_t ksteps,
                                             uint32_t accumulate_first) {
  const uint32_t idesc = make_idesc(128, N, 0, b_mn);
  for (uint32_t ks = 0; ks < ksteps; ++ks) {
    uint64_t bd = b_mn ? desc_mnmajor(b_saddr, b_rows, ks, b_col0) : desc_kmajor(b_saddr, b_rows, ks);
    mma_f16_ts(tmem_d, tmem_a + ks * 8, bd, idesc, (ks > 0) ? 1u : accumulate_first);
  }
}

}  // namespace tc05
