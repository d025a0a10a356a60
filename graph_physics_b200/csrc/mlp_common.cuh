// mlp_common.cuh -- building blocks shared by the fused MeshGraphNet MLP kernels.
//
// Work decomposition used by every kernel here: a CTA holds the block's fp16 weight
// tiles in shared memory (loaded once by bulk async copy) and runs one or two
// independent "workgroups" of 128 threads; a workgroup owns one 128-row tile at a
// time (rows = edges or nodes), its own operand buffers, its own TMEM columns and
// its own mbarrier.  Thread t of a workgroup owns row t of the accumulator
// (TMEM lane t) in the per-layer epilogues; global traffic uses a second,
// coalesced mapping (half-warp or warp per 512-byte row).
#pragma once
#include "common.cuh"
#include "tc05.cuh"

namespace mgn {
using namespace tc05;

static constexpr int H = 128;            // hidden size the kernels are built for
static constexpr int TILE_ROWS = 128;
static constexpr uint32_t TILE_BYTES = 32768;   // 128 x 128 fp16 panel tile
static constexpr float RMS_EPS = 1e-8f;
static constexpr float INV_SQRT_H = 0.08838834764831845f;  // 128^-1/2

__device__ __forceinline__ float4 ldg_f4(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }
__device__ __forceinline__ float4 ld_stream_f4(const float* p) {
  float4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ void st_stream_f4(float* p, float4 v) {
  asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z),
               "f"(v.w) : "memory");
}
__device__ __forceinline__ void red_add_f4(float* p, float4 v) {
  asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}
__device__ __forceinline__ uint4 ldg_u4(const void* p) { return __ldg(reinterpret_cast<const uint4*>(p)); }

__device__ __forceinline__ void unpack8(uint4 u, float* f) {
  const __half2* h = reinterpret_cast<const __half2*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 t = __half22float2(h[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}

// Load 32 rows [row0, row0+32) of a 128-row tile: fp32 global rows (512 B each, one
// warp instruction per row, coalesced) -> fp16 panel tile columns [col0, col0+128).
// `src_row(r)` gives the global row index or -1 (zero fill).  `mul` scales before the
// conversion (gradient scale; 1 for activations).
template <class RowFn>
__device__ __forceinline__ void load_rows_f32_to_tile(uint8_t* tile, uint32_t tile_rows, uint32_t col0,
                                                      const float* __restrict__ src, int row0, int lane,
                                                      float mul, RowFn src_row) {
#pragma unroll 1
  for (int rb = 0; rb < 32; rb += 8) {
    float4 v[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      long long g = src_row(row0 + rb + i);
      v[i] = (g >= 0) ? ld_stream_f4(src + g * H + 4 * lane) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      uint2 pk;
      pk.x = pack_h2(v[i].x * mul, v[i].y * mul);
      pk.y = pack_h2(v[i].z * mul, v[i].w * mul);
      *reinterpret_cast<uint2*>(tile + tile_off(tile_rows, row0 + rb + i, col0 + 4 * lane)) = pk;
    }
  }
}

// Copy 32 rows of an fp16 [*,128] global matrix (256 B rows) into a panel tile:
// half-warp per row, 16 B per lane.
template <class RowFn>
__device__ __forceinline__ void load_rows_f16_to_tile(uint8_t* tile, uint32_t tile_rows, uint32_t col0,
                                                      const __half* __restrict__ src, int row0, int lane,
                                                      RowFn src_row) {
  const int hw = lane >> 4, q = lane & 15;
#pragma unroll 1
  for (int rb = 0; rb < 32; rb += 16) {
    uint4 v[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      long long g = src_row(row0 + rb + 2 * i + hw);
      v[i] = (g >= 0) ? ldg_u4(src + g * H + 8 * q) : make_uint4(0, 0, 0, 0);
    }
#pragma unroll
    for (int i = 0; i < 8; ++i)
      *reinterpret_cast<uint4*>(tile + tile_off(tile_rows, row0 + rb + 2 * i + hw, col0 + 8 * q)) = v[i];
  }
}

// Row-owner epilogue store: 32 consecutive fp16 columns [c0, c0+32) of row `row`.
__device__ __forceinline__ void store_row_chunk_f16(uint8_t* tile, uint32_t tile_rows, uint32_t row, uint32_t c0,
                                                    const float* v) {
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    uint4 pk;
    pk.x = pack_h2(v[8 * j + 0], v[8 * j + 1]);
    pk.y = pack_h2(v[8 * j + 2], v[8 * j + 3]);
    pk.z = pack_h2(v[8 * j + 4], v[8 * j + 5]);
    pk.w = pack_h2(v[8 * j + 6], v[8 * j + 7]);
    *reinterpret_cast<uint4*>(tile + tile_off(tile_rows, row, c0 + 8 * j)) = pk;
  }
}
// read back 32 fp16 columns of own row (ReLU masks in the backward kernels)
__device__ __forceinline__ void load_row_chunk_f16(const uint8_t* tile, uint32_t tile_rows, uint32_t row,
                                                   uint32_t c0, float* v) {
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    uint4 u = *reinterpret_cast<const uint4*>(tile + tile_off(tile_rows, row, c0 + 8 * j));
    unpack8(u, v + 8 * j);
  }
}

// fp32 staging tile for coalesced output: 128 rows x 128 floats (64 KB), 16-byte chunks
// XOR-swizzled with row%8 so that both the row-owner writes and the row-wise reads are
// bank-conflict free.
__device__ __forceinline__ uint32_t stage_off(uint32_t row, uint32_t col) {
  uint32_t chunk = col >> 2;
  return row * 512 + (((chunk & ~7u) | ((chunk ^ row) & 7u)) << 4) + ((col & 3) << 2);
}


// Tail of every kernel's dynamic shared memory (right after the 32 KB tile slots): the block's
// bias / scale vectors (read by every epilogue; kept on chip because the tile slots leave no L1),
// the mbarriers and the TMEM base slot.  Kernels use NO static shared memory so that the dynamic
// window starts 1024-byte aligned (checked at run time).
struct SmemTail {
  float vec[5][128];      // b1..b4, scale (kernel-specific subset)
  uint64_t bar[8];
  uint32_t tmem_slot;
  uint32_t pad[3];
};
static constexpr uint32_t TAIL_BYTES = 3072;
static_assert(sizeof(SmemTail) <= TAIL_BYTES, "SmemTail too large");

__device__ __forceinline__ void check_smem_alignment(const void* base) {
  if (smem_u32(base) & 1023u) __trap();
}
__device__ __forceinline__ void copy_vec128(float* dst, const float* __restrict__ src, int tid, int nthreads) {
  if (src != nullptr)
    for (int i = tid; i < 128; i += nthreads) dst[i] = __ldg(src + i);
}

// ------------------------------------------------------------------ L2 eviction policies
// Streams that are read or written once (edge latents, gradients) are tagged evict_first so that
// they do not push the per-node arrays that every edge gathers (P, aggr, g_aggr, gP) out of L2.
__device__ __forceinline__ uint64_t policy_evict_first() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ uint64_t policy_evict_last() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ float4 ld_f4_hint(const float* p, uint64_t pol) {
  float4 r;
  asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.f32 {%0,%1,%2,%3}, [%4], %5;"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p), "l"(pol));
  return r;
}
__device__ __forceinline__ uint4 ld_u4_hint(const void* p, uint64_t pol) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.u32 {%0,%1,%2,%3}, [%4], %5;"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p), "l"(pol));
  return r;
}
__device__ __forceinline__ void st_f4_hint(float* p, float4 v, uint64_t pol) {
  asm volatile("st.global.L1::no_allocate.L2::cache_hint.v4.f32 [%0], {%1,%2,%3,%4}, %5;" ::"l"(p), "f"(v.x), "f"(v.y),
               "f"(v.z), "f"(v.w), "l"(pol) : "memory");
}
__device__ __forceinline__ void st_u4_hint(void* p, uint4 v, uint64_t pol) {
  asm volatile("st.global.L1::no_allocate.L2::cache_hint.v4.u32 [%0], {%1,%2,%3,%4}, %5;" ::"l"(p), "r"(v.x), "r"(v.y),
               "r"(v.z), "r"(v.w), "l"(pol) : "memory");
}
__device__ __forceinline__ void red_add_f4_hint(float* p, float4 v, uint64_t pol) {
  asm volatile("red.global.add.L2::cache_hint.v4.f32 [%0], {%1,%2,%3,%4}, %5;" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z),
               "f"(v.w), "l"(pol) : "memory");
}

// 256-bit global accesses (sm_100): one lane moves a whole 32-byte sector, so a thread that owns 32 consecutive
// floats of a row (the TMEM epilogue mapping) reads / writes full sectors without a shared-memory transpose
struct f8 { float v[8]; };
__device__ __forceinline__ f8 ld_f8_hint(const float* p, uint64_t pol) {
  f8 r;
  asm volatile("ld.global.L1::no_allocate.L2::cache_hint.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8], %9;"
               : "=f"(r.v[0]), "=f"(r.v[1]), "=f"(r.v[2]), "=f"(r.v[3]), "=f"(r.v[4]), "=f"(r.v[5]), "=f"(r.v[6]),
                 "=f"(r.v[7]) : "l"(p), "l"(pol));
  return r;
}
__device__ __forceinline__ void st_f8_hint(float* p, const float* v, uint64_t pol) {
  asm volatile("st.global.L1::no_allocate.L2::cache_hint.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8}, %9;" ::"l"(p), "f"(v[0]),
               "f"(v[1]), "f"(v[2]), "f"(v[3]), "f"(v[4]), "f"(v[5]), "f"(v[6]), "f"(v[7]), "l"(pol) : "memory");
}
// 16-byte asynchronous global -> shared copy (LDGSTS); src_bytes = 0 zero-fills the destination
__device__ __forceinline__ void cp_async16(uint32_t dst_saddr, const void* src, uint32_t src_bytes, uint64_t pol) {
  asm volatile("cp.async.cg.shared.global.L2::cache_hint [%0], [%1], 16, %2, %3;" ::"r"(dst_saddr), "l"(src),
               "r"(src_bytes), "l"(pol) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// ------------------------------------------------------------------ bulk (TMA) shared -> global
// Asynchronous, issued by one thread, no LSU / L1 involvement.  Completion is tracked per issuing thread in
// bulk async-groups: commit, then wait_read (source may be overwritten) or wait_all (writes performed).
__device__ __forceinline__ void bulk_s2g(void* gdst, const void* smem_src, uint32_t bytes, uint64_t pol) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group.L2::cache_hint [%0], [%1], %2, %3;" ::"l"(gdst),
               "r"(smem_u32(smem_src)), "r"(bytes), "l"(pol) : "memory");
}
// global[i] += shared[i] (fp32), performed at L2
__device__ __forceinline__ void bulk_red_add_f32(float* gdst, const void* smem_src, uint32_t bytes, uint64_t pol) {
  asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.L2::cache_hint.add.f32 [%0], [%1], %2, %3;" ::"l"(gdst),
               "r"(smem_u32(smem_src)), "r"(bytes), "l"(pol) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read1() { asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait0() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

// ------------------------------------------------------------------ developer timeline (tools/timeline.py)
// When a buffer is registered (mgn_debug_set_timeline; host globals below), thread 0 of block 0 records clock64()
// at the phase boundaries of its tiles.  Null (the default) costs one predicated load per stamp.
#ifdef MGN_TIMELINE      // developer builds only: MGN_NVCC_EXTRA=-DMGN_TIMELINE python -m graph_physics_b200.build --force
struct Timeline {
  long long* buf;
  int n, cap;
  __device__ __forceinline__ void init(bool on, long long* b, int c) {
    buf = on ? b : nullptr;
    cap = on ? c : 0;
    n = 0;
  }
  __device__ __forceinline__ void stamp() {
    if (buf != nullptr && n < cap) buf[n++] = clock64();
  }
};
#else                    // production: the stamps compile to nothing (they cost 2.6 % of edge_bwd's stall samples,
struct Timeline {        // profiles/r02_ncu_kernels.md)
  __device__ __forceinline__ void init(bool, long long*, int) {}
  __device__ __forceinline__ void stamp() {}
};
#endif

// Developer builds with -DMGN_CTA_TRACE: thread 0 of EVERY CTA records %globaltimer at four points of the kernel into
// the timeline buffer, at [(slot * 148 + blockIdx.x) * 4 + point] (tools/cta_trace.py: when do a chained consumer's
// CTAs start relative to its producer's end?).  Compiles to nothing otherwise.
#ifdef MGN_CTA_TRACE
__device__ __forceinline__ void cta_trace(long long* buf, int cap, int slot, int point) {
  if (buf != nullptr && threadIdx.x == 0) {
    const int i = (slot * 148 + (int)blockIdx.x) * 4 + point;
    if (i < cap) {
      unsigned long long t;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
      buf[i] = (long long)t;
    }
  }
}
#else
__device__ __forceinline__ void cta_trace(long long*, int, int, int) {}
#endif

// ------------------------------------------------------------------ prefetch / misc
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
// one instruction, one thread: a contiguous range (16-byte aligned, multiple of 16 bytes) into L2
__device__ __forceinline__ void bulk_prefetch_l2(const void* p, uint32_t bytes) {
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p), "r"(bytes) : "memory");
}
// ------------------------------------------------------------------ tile-level chaining of two launches
// A persistent kernel that ends in a wave tail (481 node tiles on 148 SMs: 37 CTAs run a fourth tile while 111 SMs
// idle) can hand its tiles to the NEXT launch one by one: the producer publishes ready[tile] = 1 after the tile's
// last store, the consumer -- launched with programmatic stream serialization, so its CTAs start on the SMs the
// producer's CTAs leave -- takes tiles from a counter in index order (= the order the producer finishes them in),
// waits for ready[tile], clears it for the next launch, and calls griddepcontrol.wait only when its tiles are done
// (so that ITS successor still sees "predecessor complete => everything before it complete").  Every producer CTA is
// resident before a consumer CTA can start, so the wait cannot deadlock; it is bounded like the mbarrier waits.
// sched[0] = next tile, sched[1] = CTAs that finished; the last one to finish zeroes both.
__device__ __forceinline__ void tile_publish(int32_t* ready, long long tile) {      // one thread, after a CTA barrier
  __threadfence();
  asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(ready + tile), "r"(1) : "memory");
}
__device__ __forceinline__ void tile_acquire_tagged(int32_t* ready, long long tile, uint32_t tag) {   // one thread
  int v;
  unsigned long long t0 = 0;
#pragma unroll 1
  for (uint32_t spin = 0;; ++spin) {
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(ready + tile) : "memory");
    if (v != 0) break;
    if (spin >= MBAR_FAST_POLLS) {
      unsigned long long t;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
      if (spin == MBAR_FAST_POLLS) t0 = t;
      else if (t - t0 > MBAR_TIMEOUT_NS) mbar_timeout(0xf1a9u /* a ready flag, not a barrier */, (uint32_t)tile, tag);
      if (spin == 0xffffffffu) spin = MBAR_FAST_POLLS;
    }
  }
  ready[tile] = 0;
}
#define tile_acquire(ready, tile) \
  tile_acquire_tagged((ready), (tile), ((uint32_t)(MGN_FILE_ID) << 16) | (uint32_t)__LINE__)
__device__ __forceinline__ void sched_done(int32_t* sched) {                        // one thread, at the CTA's end
  if (atomicAdd(sched + 1, 1) == (int)gridDim.x - 1) {
    sched[0] = 0;
    __threadfence();
    sched[1] = 0;
  }
}

// one thread: the rows [row0, row0 + 128) of a row-major array (row_bytes a multiple of 16) into L2 -- the persistent
// kernels call it for the tile they will process NEXT, a whole tile time ahead of its loads
__device__ __forceinline__ void prefetch_tile_rows_l2(const void* base, long long row0, long long rows_total,
                                                      uint32_t row_bytes) {
  if (row0 >= rows_total) return;
  const long long nr = (rows_total - row0) < TILE_ROWS ? (rows_total - row0) : TILE_ROWS;
  bulk_prefetch_l2(reinterpret_cast<const uint8_t*>(base) + row0 * row_bytes, (uint32_t)(nr * row_bytes));
}
// one 512-byte fp32 row = 4 lines
__device__ __forceinline__ void prefetch_row_f32(const float* row) {
  const char* c = reinterpret_cast<const char*>(row);
  prefetch_l2(c); prefetch_l2(c + 128); prefetch_l2(c + 256); prefetch_l2(c + 384);
}
__device__ __forceinline__ void prefetch_row_f16(const __half* row) {   // 256 bytes
  const char* c = reinterpret_cast<const char*>(row);
  prefetch_l2(c); prefetch_l2(c + 128);
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ float4 ld_f4(const float* p) { return *reinterpret_cast<const float4*>(p); }

// Column sums over the 32 lanes of a warp: lane j returns sum over lanes of v[j].
// Destroys v.  (butterfly transpose-reduce: 31 shuffles)
__device__ __forceinline__ float warp_colsum32(float (&v)[32], int lane) {
#pragma unroll
  for (int s = 16; s >= 1; s >>= 1) {
    const bool up = (lane & s) != 0;
#pragma unroll
    for (int i = 0; i < s; ++i) {
      const float send = up ? v[i] : v[i + s];
      const float keep = up ? v[i + s] : v[i];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, s);
    }
  }
  return v[0];
}

// Column sums of a 128 x 128 fp16 panel tile over all 128 rows: thread t (0..127) adds
// rows [64*(t>>6), +64) of columns 2*(t&63), 2*(t&63)+1 into a0, a1.
__device__ __forceinline__ void tile_colsum_pair(const uint8_t* tile, int t, float& a0, float& a1) {
  const int cp = t & 63, rbase = (t >> 6) * 64;
#pragma unroll 8
  for (int r = 0; r < 64; ++r) {
    const uint32_t u = *reinterpret_cast<const uint32_t*>(tile + tile_off(128, rbase + r, 2 * cp));
    const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&u));
    a0 += f.x;
    a1 += f.y;
  }
}

// Load NR rows (NR <= 32, multiple of BATCH) of fp32 [*,128] -> fp16 panel tile, BATCH rows in flight.
template <int BATCH, class RowFn>
__device__ __forceinline__ void load_rows_f32_to_tile_b(uint8_t* tile, uint32_t tile_rows, uint32_t col0,
                                                        const float* __restrict__ src, int row0, int nrows,
                                                        int lane, RowFn src_row) {
#pragma unroll 1
  for (int rb = 0; rb < nrows; rb += BATCH) {
    float4 v[BATCH];
#pragma unroll
    for (int i = 0; i < BATCH; ++i) {
      long long g = src_row(row0 + rb + i);
      v[i] = (g >= 0) ? ld_stream_f4(src + g * H + 4 * lane) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int i = 0; i < BATCH; ++i) {
      uint2 pk;
      pk.x = pack_h2(v[i].x, v[i].y);
      pk.y = pack_h2(v[i].z, v[i].w);
      *reinterpret_cast<uint2*>(tile + tile_off(tile_rows, row0 + rb + i, col0 + 4 * lane)) = pk;
    }
  }
}

// tile row <- fp16(a[ra(r)] + b[rb(r)]) ; either source may be absent (row fn returns -1)
template <int BATCH, class RowFnA, class RowFnB>
__device__ __forceinline__ void load_rows_sum2_to_tile(uint8_t* tile, uint32_t col0, const float* __restrict__ a,
                                                       const float* __restrict__ b, int row0, int nrows, int lane,
                                                       RowFnA ra, RowFnB rb_) {
#pragma unroll 1
  for (int rb = 0; rb < nrows; rb += BATCH) {
    float4 va[BATCH], vb[BATCH];
#pragma unroll
    for (int i = 0; i < BATCH; ++i) {
      long long ga = ra(row0 + rb + i), gb = rb_(row0 + rb + i);
      va[i] = (ga >= 0) ? ld_stream_f4(a + ga * H + 4 * lane) : make_float4(0.f, 0.f, 0.f, 0.f);
      vb[i] = (gb >= 0) ? ld_stream_f4(b + gb * H + 4 * lane) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int i = 0; i < BATCH; ++i) {
      uint2 pk;
      pk.x = pack_h2(va[i].x + vb[i].x, va[i].y + vb[i].y);
      pk.y = pack_h2(va[i].z + vb[i].z, va[i].w + vb[i].w);
      *reinterpret_cast<uint2*>(tile + tile_off(128, row0 + rb + i, col0 + 4 * lane)) = pk;
    }
  }
}

// ------------------------------------------------------------------ shared epilogues
// hidden layer: tile row t <- fp16(relu(acc + bias))
__device__ __forceinline__ void epi_bias_relu(uint32_t tmem_row, const float* __restrict__ bias, uint8_t* tile,
                                              int t) {
#pragma unroll 1
  for (int c0 = 0; c0 < H; c0 += 32) {
    float v[32];
    tmem_ld32(tmem_row + c0, v);
    float4 bb[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) bb[j] = ld_f4(bias + c0 + 4 * j);
    tmem_ld_wait();
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      v[4 * j + 0] = fmaxf(v[4 * j + 0] + bb[j].x, 0.f);
      v[4 * j + 1] = fmaxf(v[4 * j + 1] + bb[j].y, 0.f);
      v[4 * j + 2] = fmaxf(v[4 * j + 2] + bb[j].z, 0.f);
      v[4 * j + 3] = fmaxf(v[4 * j + 3] + bb[j].w, 0.f);
    }
    store_row_chunk_f16(tile, 128, t, c0, v);
  }
}
// ReLU backward: tile row t holds h = relu(z); overwrite with fp16(acc * [h > 0])
__device__ __forceinline__ void epi_relu_mask_inplace(uint32_t tmem_row, uint8_t* tile, int t) {
#pragma unroll 1
  for (int c0 = 0; c0 < H; c0 += 32) {
    float v[32], h[32];
    tmem_ld32(tmem_row + c0, v);
    load_row_chunk_f16(tile, 128, t, c0, h);
    tmem_ld_wait();
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = h[j] > 0.f ? v[j] : 0.f;
    store_row_chunk_f16(tile, 128, t, c0, v);
  }
}
// epi_bias_relu + optional fp16 copy of the row chunk to save[grow, :]
__device__ __forceinline__ void epi_bias_relu_save(uint32_t tmem_row, const float* __restrict__ bias, uint8_t* tile,
                                                   int t, __half* save, long long grow) {
#pragma unroll 1
  for (int c0 = 0; c0 < H; c0 += 32) {
    float v[32];
    tmem_ld32(tmem_row + c0, v);
    float4 bb[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) bb[j] = ld_f4(bias + c0 + 4 * j);
    tmem_ld_wait();
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      v[4 * j + 0] = fmaxf(v[4 * j + 0] + bb[j].x, 0.f);
      v[4 * j + 1] = fmaxf(v[4 * j + 1] + bb[j].y, 0.f);
      v[4 * j + 2] = fmaxf(v[4 * j + 2] + bb[j].z, 0.f);
      v[4 * j + 3] = fmaxf(v[4 * j + 3] + bb[j].w, 0.f);
    }
    store_row_chunk_f16(tile, 128, t, c0, v);
    if (save != nullptr) {
#pragma unroll
      for (int j = 0; j < 4; ++j)
        *reinterpret_cast<uint4*>(save + grow * H + c0 + 8 * j) =
            *reinterpret_cast<const uint4*>(tile + tile_off(128, t, c0 + 8 * j));
    }
  }
}
// RMSNorm backward (layers.py:49-74 differentiated): acc = z - bias, gtile row t = upstream
// gradient g of y = scale * z * inv, inv = 1 / (||z|| d^-1/2 + eps).  Overwrites gtile row t with
// fp16(dL/dz) and adds this warp's column sums of g * z * inv (= dL/dscale) to dsc[0..3]
// (lane j of the warp holds column 32*k + j in dsc[k]).
__device__ __forceinline__ void epi_rmsnorm_bwd(uint32_t tmem_row, const float* __restrict__ bias,
                                                const float* __restrict__ scale, uint8_t* gtile, int t, int lane,
                                                float (&dsc)[4]) {
  float ss = 0.f, dot = 0.f;
#pragma unroll 1
  for (int c0 = 0; c0 < H; c0 += 32) {
    float v[32], g[32];
    tmem_ld32(tmem_row + c0, v);
    load_row_chunk_f16(gtile, 128, t, c0, g);
    float4 bb[8], sc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { bb[j] = ld_f4(bias + c0 + 4 * j); sc[j] = ld_f4(scale + c0 + 4 * j); }
    tmem_ld_wait();
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float z0 = v[4 * j] + bb[j].x, z1 = v[4 * j + 1] + bb[j].y, z2 = v[4 * j + 2] + bb[j].z,
                  z3 = v[4 * j + 3] + bb[j].w;
      ss += z0 * z0 + z1 * z1 + z2 * z2 + z3 * z3;
      dot += sc[j].x * g[4 * j] * z0 + sc[j].y * g[4 * j + 1] * z1 + sc[j].z * g[4 * j + 2] * z2 +
             sc[j].w * g[4 * j + 3] * z3;
    }
  }
  const float r = sqrtf(ss) * INV_SQRT_H;
  const float inv = 1.0f / (r + RMS_EPS);
  const float k = r > 0.f ? dot * inv * inv / ((float)H * r) : 0.f;
#pragma unroll
  for (int cq = 0; cq < 4; ++cq) {
    const int c0 = cq * 32;
    float v[32], g[32], q[32];
    tmem_ld32(tmem_row + c0, v);
    load_row_chunk_f16(gtile, 128, t, c0, g);
    float4 bb[8], sc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { bb[j] = ld_f4(bias + c0 + 4 * j); sc[j] = ld_f4(scale + c0 + 4 * j); }
    tmem_ld_wait();
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float b_[4] = {bb[j].x, bb[j].y, bb[j].z, bb[j].w};
      const float s_[4] = {sc[j].x, sc[j].y, sc[j].z, sc[j].w};
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float z = v[4 * j + i] + b_[i];
        const float gg = g[4 * j + i];
        q[4 * j + i] = gg * z * inv;
        v[4 * j + i] = s_[i] * gg * inv - z * k;
      }
    }
    store_row_chunk_f16(gtile, 128, t, c0, v);
    dsc[cq] += warp_colsum32(q, lane);
  }
}
// fp32 accumulator rows -> global with red.add (thread t owns row t): dst row pitch `ld` floats
__device__ __forceinline__ void flush_acc_rows(uint32_t tmem_row, float* __restrict__ dst, int ld, int t) {
#pragma unroll 1
  for (int c0 = 0; c0 < H; c0 += 32) {
    float v[32];
    tmem_ld32(tmem_row + c0, v);
    tmem_ld_wait();
#pragma unroll
    for (int j = 0; j < 8; ++j)
      red_add_f4(dst + (long long)t * ld + c0 + 4 * j, make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]));
  }
}
// accumulator rows -> fp32 staging tile (stage_off layout, 64 KB)
__device__ __forceinline__ void acc_to_stage(uint32_t tmem_row, uint8_t* stage, int t) {
#pragma unroll 1
  for (int c0 = 0; c0 < H; c0 += 32) {
    float v[32];
    tmem_ld32(tmem_row + c0, v);
    tmem_ld_wait();
#pragma unroll
    for (int j = 0; j < 8; ++j)
      *reinterpret_cast<float4*>(stage + stage_off(t, c0 + 4 * j)) =
          make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
  }
}

}  // namespace mgn
