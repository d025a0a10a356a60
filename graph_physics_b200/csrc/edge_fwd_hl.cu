// edge_fwd_hl.cu -- the edge forward kernel of steps 1..K-1 with the residual stream kept as TWO fp16 tile images:
// hi = fp16(e) (the layer-1 operand tile, which the edge backward also recomputes from) and lo = fp16(e - hi).
//
// Same arithmetic as mlp_fwd_kernel<true> in mlp_fwd.cu (GraphNetBlock.forward, graphphysics/models/layers.py:687-698:
// edge_block on [e | x[col] | x[row]], RMSNorm, residual, sum at the receivers), except that the residual add reads
// e = hi + lo (exact in fp32, >= 21 significant bits of the fp32 stream) instead of the fp32 rows.  What this buys:
//   * the edge stream moves ONLY by bulk copies -- hi / lo tiles in, hi / lo tiles out, 4 x 32 KB per 128 edges instead
//     of 2 x 64 KB fp32 rows + 2 x 32 KB images -- so no thread waits on a load of the stream or drains its stores;
//   * everything after the last product runs in the accumulator's own (row, 32-column) mapping: no fp32 staging tile,
//     no second pass over the rows; the segmented sum by receiver is a shuffle scan over the 32 rows of a warp
//     (receivers are sorted, so a segment is a run of lanes) followed by one red.add per (run, column quad);
//   * the layer-1 product of the next tile is issued before that scan, so the tensor pipe works under it.
// Shared memory: four weight tiles (resident), hi tile (free after layer 1: the next tile's image is requested three
// products ahead), lo tile (updated in place), one output tile for the new hi (between tiles it stages the gathered
// P_j rows of the next tile).  Tensor memory: accumulator (128
// columns), the hidden activations as the next product's A operand (64), the parked hi chunk of every thread (64).
#define MGN_FILE_ID 2
#include <algorithm>

#include <cuda.h>                 // CUtensorMap (types only: cuTensorMapEncodeTiled is looked up at run time)
#include <cudaTypedefs.h>

#include "mlp_v2.cuh"
#include "mlp_fwd_abi.cuh"

using namespace mgn;

struct EdgeHLParams {
  long long E;
  const __half* hi_in;      // tile images of fp16(e_in)
  const __half* lo_in;      // tile images of fp16(e_in - hi)
  __half* hi_out;           // tile images of e_out, or null
  __half* lo_out;           // or null
  float* e_out;             // [E,128] fp32, CSR order, or null (only the last step needs the rows)
  const __half* P;          // [N,256] fp16
  const int32_t* senders;
  const int32_t* receivers;
  float* aggr;              // [N,128] fp32, zero on entry
  float* carry;             // [ceil(E/32),128] fp32: partial sums of runs that continue from an earlier 32-row slot
                            // (deterministic aggregation, aggr_fixup_kernel below)
  float inv_sqrt_d;         // RMSNorm: d^-1/2 of the TRUE hidden size (zero-padded narrower models, ops.py)
  int pj_tma;               // 1: the P_j[sender] rows are gathered by the TMA engine (cp.async.bulk.tensor ... tile::gather4
                            // through `tmap_p`) instead of cp.async
  int num_nodes;
  alignas(64) CUtensorMap tmap_p;   // P as a [N][256] fp16 tensor, box 64 x 1, 128-byte swizzle
  __half* h_save[3];        // training: tile images of the hidden activations h1, h2, h3 (what mgn_edge_bwd_saved starts
                            // from instead of recomputing them), all three or all null
  const __half* W;          // We, W2, W3, W4 tiles
  const float* b2;
  const float* b3;
  const float* b4;
  const float* scale;
  long long* dbg;
  int dbg_cap;
};

// SAVE (training): the hidden activations h1, h2, h3 also leave as tile images.  They are staged in shared-memory tiles
// that are idle at that point -- h1 in sHi (its hi chunks are parked in tensor memory), h2 in sLo (the remainder tile
// is only needed after the last product), h3 in sHi again -- so that every 32 KB copy has two layer times to read its
// tile before the tile is written again; the loads of lo (this tile) and hi (next tile) follow those copies, and the
// next tile's first product is issued after the scan instead of before it.
// four rows of a [N][256] fp16 tensor (64 columns from `col0`) -> four consecutive 128-byte rows of a swizzled panel
__device__ __forceinline__ void tma_gather4(void* smem_dst, const CUtensorMap* tmap, int col0, int r0, int r1, int r2,
                                            int r3, uint64_t* bar, uint64_t pol) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cta.global.tile::gather4.mbarrier::complete_tx::bytes.L2::cache_hint"
      " [%0], [%1, {%2, %3, %4, %5, %6}], [%7], %8;" ::"r"(smem_u32(smem_dst)),
      "l"(tmap), "r"(col0), "r"(r0), "r"(r1), "r"(r2), "r"(r3), "r"(smem_u32(bar)), "l"(pol)
      : "memory");
}

template <bool SAVE, bool PJ_TMA>
__global__ void __launch_bounds__(NT, 1) edge_fwd_hl_kernel(const __grid_constant__ EdgeHLParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  check_smem_alignment(smem);
  uint8_t* sW = smem;                          // 4 x 32 KB
  uint8_t* sHi = smem + 4 * TILE_BYTES;
  uint8_t* sLo = smem + 5 * TILE_BYTES;
  uint8_t* sOut = smem + 6 * TILE_BYTES;
  Tail2* tail = reinterpret_cast<Tail2*>(smem + 7 * TILE_BYTES);
  uint64_t& wbar = tail->bar[0];
  uint64_t& mbar = tail->bar[1];
  uint64_t& bar_hi = tail->bar[2];
  uint64_t& bar_lo = tail->bar[3];
  uint64_t& bar_pj = tail->bar[4];
  uint32_t& tmem_slot = tail->tmem_slot;
  const Map m = make_map();

  if (m.tid == 0) {
    mbar_init(&wbar, 1);
    mbar_init(&mbar, 1);
    mbar_init(&bar_hi, 1);
    mbar_init(&bar_lo, 1);
    mbar_init(&bar_pj, 1);
    mbar_init_fence();
  }
  if (m.warp == 0) tmem_alloc<512>(&tmem_slot);
  // the first tile's image and remainder (written by the previous step's edge kernel, three launches back -- not by
  // the kernel griddepcontrol.wait is for) into L2 while the previous kernel drains
  if (m.tid == 64 && (long long)blockIdx.x * TILE_ROWS < p.E) {
    bulk_prefetch_l2(reinterpret_cast<const uint8_t*>(p.hi_in) + (size_t)blockIdx.x * TILE_BYTES, TILE_BYTES);
    bulk_prefetch_l2(reinterpret_cast<const uint8_t*>(p.lo_in) + (size_t)blockIdx.x * TILE_BYTES, TILE_BYTES);
  }
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  pdl_wait();        // the previous kernel in the stream is complete: global memory may be touched from here on
  pdl_launch();      // (common.cuh) the next kernel may start its own prologue under this one
  const uint32_t tmem_acc = tmem_slot;
  const uint32_t taddr = tmem_acc + m.lane_off + m.c0;
  const uint32_t tmem_a = tmem_acc + 128;
  const uint32_t taddr_a = tmem_a + m.lane_off + (m.c0 >> 1);
  const uint32_t taddr_park = tmem_acc + 192 + m.lane_off + (m.c0 >> 1);
  // row sums of the RMSNorm across the four column quarters go through four tensor-memory columns (one per quarter,
  // lanes = rows) instead of shared memory; the 2 KB this frees hold b2, b3, b4 and scale, so no epilogue waits for a
  // global load of its bias behind the gathers in the load/store queue
  const uint32_t taddr_rs = tmem_acc + 256 + m.lane_off;
  float* svec = tail->red;                     // [4][128]: b2, b3, b4, scale
  if (m.tid < 128) {
    svec[m.tid] = __ldg(p.b2 + m.tid);
    svec[128 + m.tid] = __ldg(p.b3 + m.tid);
    svec[256 + m.tid] = __ldg(p.b4 + m.tid);
    svec[384 + m.tid] = __ldg(p.scale + m.tid);
  }
  __syncthreads();
  auto vec32 = [&](int which, float4 (&b)[8]) {
#pragma unroll
    for (int j = 0; j < 8; ++j) b[j] = *reinterpret_cast<const float4*>(svec + which * 128 + m.c0 + 4 * j);
  };
  const uint32_t sW_u32 = smem_u32(sW), sHi_u32 = smem_u32(sHi);

  const long long num_tiles = (p.E + TILE_ROWS - 1) / TILE_ROWS;
  const long long first = blockIdx.x, stride = gridDim.x;
  const int K = first < num_tiles ? (int)((num_tiles - first + stride - 1) / stride) : 0;
  auto tile_of = [&](int k) -> long long { return first + (long long)k * stride; };
  const uint64_t pol_stream = policy_evict_first();
  const uint64_t pol_keep = policy_evict_last();

  if (m.tid == 0) {
    mbar_arrive_expect_tx(&wbar, 4 * TILE_BYTES);
    for (uint32_t o = 0; o < 4 * TILE_BYTES; o += TILE_BYTES)
      bulk_g2s(sW + o, reinterpret_cast<const uint8_t*>(p.W) + o, TILE_BYTES, &wbar);
  }
  auto issue_hi = [&](int k) {                 // thread 0
    mbar_arrive_expect_tx(&bar_hi, TILE_BYTES);
    bulk_g2s(sHi, reinterpret_cast<const uint8_t*>(p.hi_in) + (size_t)tile_of(k) * TILE_BYTES, TILE_BYTES, &bar_hi);
  };
  auto issue_lo = [&](int k) {                 // thread 0
    mbar_arrive_expect_tx(&bar_lo, TILE_BYTES);
    bulk_g2s(sLo, reinterpret_cast<const uint8_t*>(p.lo_in) + (size_t)tile_of(k) * TILE_BYTES, TILE_BYTES, &bar_lo);
  };
  uint32_t phase = 0;
  auto mma_wait = [&]() {
    mbar_wait(&mbar, phase);
    phase ^= 1;
    fence_after_sync();
  };
  constexpr bool save = SAVE;
  auto acts_ready = [&]() {       // the TMEM-resident activations (and parked rows) of every warp are written
    tmem_st_wait();
    if (save) fence_async_smem();               // ... and the copy of them in sHi is visible to the bulk-copy engine
    fence_before_sync();
    __syncthreads();
  };
  // training: the hidden activation chunk of this thread (32 fp32 -> 16 packed words) also goes to its place in sHi
  auto save_chunk = [&](uint8_t* dst, const float (&v)[32]) {
#pragma unroll
    for (int j = 0; j < 4; ++j)
      *reinterpret_cast<uint4*>(dst + tile_off(128, m.row, m.c0 + 8 * j)) =
          make_uint4(pack_h2(v[8 * j], v[8 * j + 1]), pack_h2(v[8 * j + 2], v[8 * j + 3]),
                     pack_h2(v[8 * j + 4], v[8 * j + 5]), pack_h2(v[8 * j + 6], v[8 * j + 7]));
  };
  // thread 0: staging tile -> image `which` of tile `tile_` (its own bulk group)
  auto store_save = [&](int which, long long tile_, const uint8_t* src) {
    bulk_s2g(reinterpret_cast<uint8_t*>(p.h_save[which]) + (size_t)tile_ * TILE_BYTES, src, TILE_BYTES, pol_stream);
    bulk_commit();
  };
  // indices, requested one tile ahead: this thread's row (epilogue mapping) -> receiver; coalesced mapping (lane j < 8
  // of warp w stands for tile row 8w + j) -> sender
  struct Idx { int rcv, snd_c, prev; bool ok; };      // prev: receiver of the row before this warp's 32-row slot (-1: none)
  auto load_idx = [&](int k) -> Idx {
    Idx ix{0, -1, -1, false};
    if (k >= K) return ix;
    const long long r0 = tile_of(k) * TILE_ROWS;
    if (r0 + m.row < p.E) {
      ix.rcv = __ldg(p.receivers + r0 + m.row);
      ix.ok = true;
    }
    const long long s0 = r0 + m.lq * 32;
    if (s0 > 0 && s0 < p.E) ix.prev = __ldg(p.receivers + s0 - 1);
    const long long g = r0 + m.warp * 8 + (m.lane & 7);
    if (g < p.E) ix.snd_c = __ldg(p.senders + g);
    return ix;
  };
  // P_i[receiver]: read directly by the row's thread (rows of one receiver are consecutive, so a warp's request
  // touches a handful of lines).  P_j[sender] is a true gather: asynchronous copies into the idle output tile, half
  // a warp per 256-byte row, picked up from there by the row's thread before the tile is needed for the new image.
  uint4 pi[4], pj[4];
  auto load_pi = [&](const Idx& ix) {
    const __half* Pi = p.P + (long long)ix.rcv * 256 + m.c0;
#pragma unroll
    for (int j = 0; j < 4; ++j) pi[j] = ix.ok ? ld_u4_hint(Pi + 8 * j, pol_keep) : make_uint4(0, 0, 0, 0);
  };
  const uint32_t sOut_u32 = smem_u32(sOut);
  uint32_t pj_phase = 0;
  auto stage_pj = [&](const Idx& ix) {
    if constexpr (PJ_TMA) {
      // TMA gather: lanes 0..3 of every warp move the warp's 8 rows -- (row group of 4) x (64-column panel) each, one
      // cp.async.bulk.tensor.2d.tile::gather4 = 4 x 128 bytes, landing as rows 4g..4g+3 of the panel exactly as
      // tile_off lays them out (the 128-byte swizzle is a function of the shared-memory address).  Rows past the end
      // name row N: out of bounds, filled with zeros.
      int sn[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        sn[i] = __shfl_sync(0xffffffffu, ix.snd_c, i);
        if (sn[i] < 0) sn[i] = p.num_nodes;
      }
      if (m.tid == 0) mbar_arrive_expect_tx(&bar_pj, TILE_BYTES);
      if (m.lane < 4) {
        const int g = m.lane & 1, pn = m.lane >> 1;
        fence_async_smem();
        tma_gather4(sOut + pn * (TILE_ROWS * 128) + (m.warp * 8 + 4 * g) * 128, &p.tmap_p, 128 + 64 * pn,
                    g ? sn[4] : sn[0], g ? sn[5] : sn[1], g ? sn[6] : sn[2], g ? sn[7] : sn[3], &bar_pj, pol_keep);
      }
      return;
    }
    const int hw = m.lane >> 4, q = m.lane & 15;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int sn = __shfl_sync(0xffffffffu, ix.snd_c, 2 * i + hw);
      cp_async16(sOut_u32 + tile_off(128, m.warp * 8 + 2 * i + hw, 8 * q),
                 p.P + (long long)(sn >= 0 ? sn : 0) * 256 + 128 + 8 * q, sn >= 0 ? 16u : 0u, pol_keep);
    }
  };
  auto wait_pj = [&]() {                       // this thread may read the staged rows
    if constexpr (PJ_TMA) {
      mbar_wait(&bar_pj, pj_phase);
      pj_phase ^= 1;
    } else {
      cp_async_wait_all();
    }
  };
  auto fetch_pj = [&]() {                      // after wait_pj() (and, for cp.async, a barrier)
#pragma unroll
    for (int j = 0; j < 4; ++j) pj[j] = *reinterpret_cast<const uint4*>(sOut + tile_off(128, m.row, m.c0 + 8 * j));
  };
  auto issue_l1 = [&](int k) {                 // thread 0: product 1 of the k-th tile of this CTA
    mbar_wait(&bar_hi, (uint32_t)k & 1u);
    fence_after_sync();
    umma_tile(tmem_acc, sHi_u32, 128, 0, sW_u32, 128, 0, 0, 128, 8, 0);
    mma_commit(&mbar);
  };
  Timeline tl;
  tl.init(m.tid == 0 && blockIdx.x == 0, p.dbg, p.dbg_cap);

  // ---- prologue: first tile's image, indices, P rows and product 1
  Idx cur = load_idx(0);
  Idx nxt = load_idx(1);
  if (K > 0) {
    if (m.tid == 0) {
      issue_hi(0);
      mbar_wait(&wbar, 0);
      issue_l1(0);
    }
    load_pi(cur);
    stage_pj(cur);
    wait_pj();
    __syncthreads();
    fetch_pj();
  }

  for (int k = 0; k < K; ++k) {
    tl.stamp();   // 0 tile start (product 1 in flight)
    const long long tile = tile_of(k);
    const long long grow = tile * TILE_ROWS + m.row;
    const bool rvalid = cur.ok;
    const bool has_next = k + 1 < K;

    // ---- layer 1: + P_i[receiver] + P_j[sender], ReLU -> A operand in tensor memory; park this thread's hi chunk
    mma_wait();
    tl.stamp();   // 1 product 1 done
    {
      float v[32];
      tmem_ld32(taddr, v);
      uint32_t hr[16];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const uint4 u = *reinterpret_cast<const uint4*>(sHi + tile_off(128, m.row, m.c0 + 8 * j));
        hr[4 * j] = u.x; hr[4 * j + 1] = u.y; hr[4 * j + 2] = u.z; hr[4 * j + 3] = u.w;
      }
      tmem_st16(taddr_park, hr);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float a[8], b[8];
        unpack8(pi[j], a);
        unpack8(pj[j], b);
#pragma unroll
        for (int i = 0; i < 8; ++i) v[8 * j + i] = fmaxf(v[8 * j + i] + a[i] + b[i], 0.f);
      }
      store_row_chunk_tmem(taddr_a, v, nullptr, 0, m.c0);
      if (save) save_chunk(sHi, v);            // this thread's hi chunk is parked: the same bytes of sHi take h1
    }
    acts_ready();
    if (m.tid == 0) {
      fence_after_sync();
      umma_tile_ts(tmem_acc, tmem_a, sW_u32 + TILE_BYTES, 128, 0, 0, 128, 8, 0);
      if (!save) {
        mma_commit(&mbar);
        if (has_next) issue_hi(k + 1);         // the hi tile is parked: next tile's operand image, three products ahead
        if (k > 0) bulk_wait_read0();          // the previous tile's stores have read sLo / sOut
        issue_lo(k);
      } else {
        if (k > 0) bulk_wait_read0();          // the previous tile's stores have read sLo (the next epilogue puts h2
        mma_commit(&mbar);                     // there) and sOut; issued a scan + an epilogue ago: no wait in practice
        store_save(0, tile, sHi);
      }
    }
    tl.stamp();   // 2 layer-1 epilogue done

    // ---- layers 2, 3
#pragma unroll 1
    for (int layer = 1; layer < 3; ++layer) {
      float4 bb[8];
      vec32(layer == 1 ? 0 : 1, bb);
      mma_wait();
      {
        float v[32];
        tmem_ld32(taddr, v);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          v[4 * j + 0] = fmaxf(v[4 * j + 0] + bb[j].x, 0.f);
          v[4 * j + 1] = fmaxf(v[4 * j + 1] + bb[j].y, 0.f);
          v[4 * j + 2] = fmaxf(v[4 * j + 2] + bb[j].z, 0.f);
          v[4 * j + 3] = fmaxf(v[4 * j + 3] + bb[j].w, 0.f);
        }
        store_row_chunk_tmem(taddr_a, v, nullptr, 0, m.c0);
        if (save) save_chunk(layer == 1 ? sLo : sHi, v);     // h2 -> sLo, h3 -> sHi
      }
      acts_ready();
      if (m.tid == 0) {
        fence_after_sync();
        umma_tile_ts(tmem_acc, tmem_a, sW_u32 + (uint32_t)(layer + 1) * TILE_BYTES, 128, 0, 0, 128, 8, 0);
        if (!save) {
          mma_commit(&mbar);
        } else if (layer == 1) {
          store_save(1, tile, sLo);
          bulk_wait_read1();                   // the copy of h1 (two layers old) has read sHi: the epilogue of the
          mma_commit(&mbar);                   // product just issued writes h3 there
        } else {
          mma_commit(&mbar);
          store_save(2, tile, sHi);
          bulk_wait_read1();                   // the copy of h2 has read sLo (this wait runs under product 4) ...
          issue_lo(k);                         // ... the remainder tile lands there for the residual
        }
      }
      // the output tile is idle (thread 0 saw the previous tile's stores leave it before this barrier): the next
      // tile's P_j rows land there under products 3 and 4
      if (layer == 1 && has_next) stage_pj(nxt);
    }
    tl.stamp();   // 3 layers 2, 3 done, product 4 issued

    // ---- layer 4 + RMSNorm; residual in image + remainder form; new images
    float y[32];
    {
      float4 bb[8];
      vec32(2, bb);
      mma_wait();
      tmem_ld32(taddr, y);
      float4 sc[8];
      vec32(3, sc);
      tmem_ld_wait();
      float ss = 0.f;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        y[4 * j] += bb[j].x; y[4 * j + 1] += bb[j].y; y[4 * j + 2] += bb[j].z; y[4 * j + 3] += bb[j].w;
        ss += y[4 * j] * y[4 * j] + y[4 * j + 1] * y[4 * j + 1] + y[4 * j + 2] * y[4 * j + 2] + y[4 * j + 3] * y[4 * j + 3];
      }
      if (has_next) wait_pj();    // (the barrier below makes every thread's staged P_j rows visible)
      tmem_st1(taddr_rs + m.cq, __float_as_uint(ss));
      tmem_st_wait();
      fence_before_sync();        // the accumulator is read: the barrier orders this before the next product 1
      __syncthreads();
      fence_after_sync();
      {
        float q4[4];
        tmem_ld4(taddr_rs, q4);
        tmem_ld_wait();
        ss = (q4[0] + q4[1]) + (q4[2] + q4[3]);
      }
      fetch_pj();                 // next tile's P_j chunk of this thread's row; the same bytes then take the new image
      const float inv = 1.0f / (sqrtf(ss) * p.inv_sqrt_d + RMS_EPS);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        y[4 * j + 0] = sc[j].x * (y[4 * j + 0] * inv);
        y[4 * j + 1] = sc[j].y * (y[4 * j + 1] * inv);
        y[4 * j + 2] = sc[j].z * (y[4 * j + 2] * inv);
        y[4 * j + 3] = sc[j].w * (y[4 * j + 3] * inv);
      }
    }
    tl.stamp();   // 4 product 4 + norm done
    {
      uint32_t hr[16];
      tmem_ld16(taddr_park, reinterpret_cast<float*>(hr));
      mbar_wait(&bar_lo, (uint32_t)k & 1u);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        uint8_t* plo = sLo + tile_off(128, m.row, m.c0 + 8 * j);
        const uint4 lu = *reinterpret_cast<const uint4*>(plo);
        const uint32_t lw[4] = {lu.x, lu.y, lu.z, lu.w};
        uint32_t hn[4], ln[4];
        float s[8];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float2 h = __half22float2(*reinterpret_cast<const __half2*>(&hr[4 * j + i]));
          const float2 l = __half22float2(*reinterpret_cast<const __half2*>(&lw[i]));
          s[2 * i] = (h.x + l.x) + y[8 * j + 2 * i];
          s[2 * i + 1] = (h.y + l.y) + y[8 * j + 2 * i + 1];
          hn[i] = pack_h2(s[2 * i], s[2 * i + 1]);
          const float2 hf = __half22float2(*reinterpret_cast<const __half2*>(&hn[i]));
          ln[i] = pack_h2(s[2 * i] - hf.x, s[2 * i + 1] - hf.y);
          if (!rvalid) { hn[i] = 0u; ln[i] = 0u; }     // rows past the end: zeros in both images
        }
        *reinterpret_cast<uint4*>(sOut + tile_off(128, m.row, m.c0 + 8 * j)) = make_uint4(hn[0], hn[1], hn[2], hn[3]);
        *reinterpret_cast<uint4*>(plo) = make_uint4(ln[0], ln[1], ln[2], ln[3]);
        if (p.e_out != nullptr && rvalid) st_f8_hint(p.e_out + grow * H + m.c0 + 8 * j, s, pol_stream);
      }
    }
    fence_async_smem();
    fence_before_sync();
    __syncthreads();
    tl.stamp();   // 5 new images staged
    if (m.tid == 0) {
      if (save) {
        bulk_wait_read0();                     // the copy of h3 (a product, the norm and the images old) has read sHi
        if (has_next) issue_hi(k + 1);
      }
      if (p.hi_out != nullptr) {
        bulk_s2g(reinterpret_cast<uint8_t*>(p.hi_out) + (size_t)tile * TILE_BYTES, sOut, TILE_BYTES, pol_stream);
        bulk_s2g(reinterpret_cast<uint8_t*>(p.lo_out) + (size_t)tile * TILE_BYTES, sLo, TILE_BYTES, pol_stream);
      }
      bulk_commit();
      if (!save && has_next) issue_l1(k + 1);  // every thread has read the accumulator: next tile's product 1
    }
    tl.stamp();   // 6 stores + next product 1 issued
    // ---- next tile's indices / P rows, then the segmented sum of y over the receivers (sorted: runs of lanes)
    const int key = rvalid ? cur.rcv : -1 - m.lane;     // rows past the end never join a run
    const int prev_rcv = cur.prev;
    const long long slot = tile * 4 + m.lq;             // this warp's 32-row slot
    cur = nxt;
    if (has_next) load_pi(cur);
    nxt = load_idx(k + 2);
    tl.stamp();   // 7 next P rows requested
    {
#pragma unroll 1
      for (int d = 1; d < 32; d <<= 1) {
        const int up = __shfl_up_sync(0xffffffffu, key, d);
        const bool same = m.lane >= d && up == key;
        if (!__any_sync(0xffffffffu, same)) break;       // no run reaches back d lanes (nor further)
#pragma unroll
        for (int c = 0; c < 32; ++c) {
          const float t = __shfl_up_sync(0xffffffffu, y[c], d);
          y[c] += same ? t : 0.f;
        }
      }
      tl.stamp();   // 8 scan done
      if (save && m.tid == 0 && has_next) issue_l1(k + 1);   // hi (requested before the scan) has landed
      const int dn = __shfl_down_sync(0xffffffffu, key, 1);
      const int key0 = __shfl_sync(0xffffffffu, key, 0);
      if (rvalid && (m.lane == 31 || dn != key)) {       // last lane of a run holds its sum over this slot
        {
          // deterministic: a run that starts in this slot owns its node's row of aggr (plain store: receivers are
          // sorted, so no other slot starts a run of that node); a run that continues from an earlier slot leaves
          // its partial in carry[slot], added to aggr in slot order by aggr_fixup_kernel after this kernel
          const bool head = key == key0 && prev_rcv == key;
          float* dst = head ? p.carry + slot * H + m.c0 : p.aggr + (long long)key * H + m.c0;
#pragma unroll
          for (int j = 0; j < 8; ++j)
            st_f4_hint(dst + 4 * j, make_float4(y[4 * j], y[4 * j + 1], y[4 * j + 2], y[4 * j + 3]), pol_keep);
        }
      }
    }
    tl.stamp();   // 9 aggregation issued
  }
  if (K == 0 && m.tid == 0) mbar_wait(&wbar, 0);    // never exit with the weight copy in flight
  if (m.tid == 0) bulk_wait0();

  fence_before_sync();
  __syncthreads();
  if (m.warp == 0) tmem_dealloc<512>(tmem_slot);
}

// ---------------------------------------------------------------------------------------------------------------
// Deterministic aggregation, second half.  The edge kernels cut the receiver-sorted edge list into slots of
// `slot_rows` rows (32: one warp's rows in edge_fwd_hl; 8: one warp's rows in mlp_fwd_kernel<true>).  A run of equal
// receivers that STARTS in a slot is stored to aggr[node] by that slot; the partial sums of its continuation in later
// slots sit in carry[slot].  Here one warp per slot finds the first continuation slot of each such run and adds the
// chain carry[s], carry[s+1], ... to aggr[node] in slot order: one writer per node, fixed order, no atomics.
struct AggrFixupParams {
  long long E, N;
  int slot_rows;
  const int32_t* receivers;
  const int32_t* rowptr;    // [N+1] receiver CSR, or null (then the caller has zeroed aggr)
  const float* carry;
  float* aggr;
};

__global__ void __launch_bounds__(256) aggr_fixup_kernel(const AggrFixupParams p) {
  pdl_wait();
  pdl_launch();
  const int lane = threadIdx.x & 31;
  const long long nslots = (p.E + p.slot_rows - 1) / p.slot_rows;
  const long long warps = ((long long)gridDim.x * blockDim.x) >> 5;
  const long long warp0 = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  // nodes nobody sends to: no edge kernel wrote their row -- zero it here (this replaces a memset of the whole array)
  if (p.rowptr != nullptr) {
    for (long long n0 = warp0 * 32; n0 < p.N; n0 += warps * 32) {
      const long long n = n0 + lane;
      const bool empty = n < p.N && __ldg(p.rowptr + n) == __ldg(p.rowptr + n + 1);
      unsigned mask = __ballot_sync(0xffffffffu, empty);
      while (mask) {
        const int b = __ffs(mask) - 1;
        mask &= mask - 1;
        reinterpret_cast<float4*>(p.aggr + (n0 + b) * H)[lane] = make_float4(0.f, 0.f, 0.f, 0.f);
      }
    }
  }
  for (long long s = warp0 + 1; s < nslots; s += warps) {
    const long long r0 = s * p.slot_rows;
    const int k = __ldg(p.receivers + r0);
    if (__ldg(p.receivers + r0 - 1) != k) continue;                       // this slot starts a new run: nothing carried
    const long long q0 = r0 - p.slot_rows;                                 // previous slot: same chain if it is a
    if (__ldg(p.receivers + q0) == k && q0 > 0 && __ldg(p.receivers + q0 - 1) == k) continue;   // continuation too
    float4 acc = __ldg(reinterpret_cast<const float4*>(p.carry + s * H) + lane);
    for (long long t = s + 1; t < nslots && __ldg(p.receivers + t * p.slot_rows) == k; ++t) {
      const float4 c = __ldg(reinterpret_cast<const float4*>(p.carry + t * H) + lane);
      acc.x += c.x; acc.y += c.y; acc.z += c.z; acc.w += c.w;
    }
    float4* dst = reinterpret_cast<float4*>(p.aggr + (long long)k * H) + lane;
    float4 o = *dst;
    o.x += acc.x; o.y += acc.y; o.z += acc.z; o.w += acc.w;
    *dst = o;
  }
}

int mgn_launch_aggr_fixup(int64_t E, int64_t N, int slot_rows, const int32_t* receivers, const int32_t* rowptr,
                          const float* carry, float* aggr, cudaStream_t stream) {
  const int64_t nslots = ceil_div64(E, slot_rows);
  if (nslots <= 1 && rowptr == nullptr) return MGN_OK;
  AggrFixupParams p{E, N, slot_rows, receivers, rowptr, carry, aggr};
  const int64_t work = std::max<int64_t>(nslots, rowptr ? ceil_div64(N, 32) : 0);
  const int grid = (int)std::min<int64_t>(std::max<int64_t>(ceil_div64(work, 8), 1), (int64_t)mgn_num_sms() * 8);
  MGN_CUDA(mgn_launch(aggr_fixup_kernel, grid, 256, 0, stream, p));
  MGN_LAUNCH_CHECK();
  return MGN_OK;
}

static constexpr int EDGE_HL_SMEM = 7 * 32768 + TAIL_BYTES;

// Tensor map of P [N][256] fp16 for the TMA gather of the P_j rows: box = 64 columns x 1 row (tile::gather4 moves four
// such boxes per instruction), 128-byte swizzle (= the panel layout of tc05.cuh), out-of-bounds rows read as zeros.
// cuTensorMapEncodeTiled is a host-side encoder of the driver library: looked up through the runtime, nothing to link.
static int make_p_tensor_map(CUtensorMap* tm, const void* P, int64_t N) {
  static PFN_cuTensorMapEncodeTiled_v12000 encode = nullptr;
  if (encode == nullptr) {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    MGN_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q));
    MGN_CHECK_ARG(fn != nullptr && q == cudaDriverEntryPointSuccess, "edge_fwd_hl: cuTensorMapEncodeTiled not available");
    encode = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(fn);
  }
  const cuuint64_t gdim[2] = {256, (cuuint64_t)N};
  const cuuint64_t gstride[1] = {512};
  const cuuint32_t box[2] = {64, 1};
  const cuuint32_t estride[2] = {1, 1};
  const CUresult r = encode(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void*>(P), gdim, gstride, box, estride,
                            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    mgn_set_error("edge_fwd_hl: cuTensorMapEncodeTiled failed (CUresult %d)", (int)r);
    return MGN_EUNSUPPORTED;
  }
  return MGN_OK;
}

// Steps 1..K-1 of the processor loop with the edge stream as image + remainder (see the header of this file).
// hi_out / lo_out (both or neither) receive the images of e_out; e_out (fp32 rows, CSR order) is optional.
// carry: [ceil(E/32),128] fp32 workspace of the deterministic aggregation (plain stores + aggr_fixup_kernel: no
// atomics).  norm_dim: the model's true hidden size (RMSNorm divides
// by sqrt(norm_dim); 128 unless the caller zero-pads a narrower model into the 128-wide kernels).
// Replaces the same reference lines as mgn_edge_fwd (graphphysics/models/layers.py:687-698).
extern "C" int mgn_edge_fwd_hl(int64_t E, int64_t N, const void* hi_in, const void* lo_in, void* hi_out, void* lo_out,
                               float* e_out, const void* P, const int32_t* senders, const int32_t* receivers,
                               float* aggr, const void* packed, const float* b2, const float* b3, const float* b4,
                               const float* scale, float* carry, const int32_t* rowptr, int norm_dim,
                               void* h1_save, void* h2_save, void* h3_save, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  MGN_CHECK_ARG(E >= 0 && N > 0 && aggr, "edge_fwd_hl: bad sizes");
  if (rowptr == nullptr || E == 0) MGN_CUDA(cudaMemsetAsync(aggr, 0, sizeof(float) * N * H, stream));
  if (E == 0) return MGN_OK;
  MGN_CHECK_ARG(hi_in && lo_in && P && senders && receivers && packed && b2 && b3 && b4 && scale && carry,
                "edge_fwd_hl: null pointer");
  MGN_CHECK_ARG((hi_out == nullptr) == (lo_out == nullptr), "edge_fwd_hl: hi_out and lo_out go together");
  MGN_CHECK_ARG(hi_out != hi_in && lo_out != lo_in, "edge_fwd_hl: in-place not supported");
  EdgeHLParams p{};
  p.E = E; p.hi_in = (const __half*)hi_in; p.lo_in = (const __half*)lo_in;
  p.hi_out = (__half*)hi_out; p.lo_out = (__half*)lo_out; p.e_out = e_out;
  p.P = (const __half*)P; p.senders = senders; p.receivers = receivers; p.aggr = aggr;
  p.W = reinterpret_cast<const __half*>(reinterpret_cast<const uint8_t*>(packed) + PK_EDGE);
  p.b2 = b2; p.b3 = b3; p.b4 = b4; p.scale = scale;
  MGN_CHECK_ARG(norm_dim > 0 && norm_dim <= H, "edge_fwd_hl: norm_dim must be in 1..128");
  p.carry = carry; p.inv_sqrt_d = 1.0f / sqrtf((float)norm_dim);
  MGN_CHECK_ARG((h1_save == nullptr) == (h2_save == nullptr) && (h1_save == nullptr) == (h3_save == nullptr),
                "edge_fwd_hl: the three hidden-activation saves go together");
  p.h_save[0] = (__half*)h1_save; p.h_save[1] = (__half*)h2_save; p.h_save[2] = (__half*)h3_save;
  // MGN_PJ_TMA=1: gather the P_j rows with the TMA engine (tile::gather4) instead of cp.async
  static const bool pj_tma = getenv("MGN_PJ_TMA") != nullptr && atoi(getenv("MGN_PJ_TMA")) != 0;
  p.pj_tma = pj_tma ? 1 : 0;
  p.num_nodes = (int)N;
  if (pj_tma) {
    MGN_CHECK_ARG((reinterpret_cast<uintptr_t>(P) & 15) == 0, "edge_fwd_hl: P must be 16-byte aligned for the tensor map");
    const int rc = make_p_tensor_map(&p.tmap_p, P, N);
    if (rc != MGN_OK) return rc;
  }
  p.dbg = g_mgn_timeline_host; p.dbg_cap = g_mgn_timeline_cap_host;
  static bool attr = false;
  if (!attr) {
    MGN_CUDA(cudaFuncSetAttribute(edge_fwd_hl_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, EDGE_HL_SMEM));
    MGN_CUDA(cudaFuncSetAttribute(edge_fwd_hl_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, EDGE_HL_SMEM));
    MGN_CUDA(cudaFuncSetAttribute(edge_fwd_hl_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, EDGE_HL_SMEM));
    MGN_CUDA(cudaFuncSetAttribute(edge_fwd_hl_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, EDGE_HL_SMEM));
    attr = true;
  }
  const int64_t tiles = ceil_div64(E, TILE_ROWS);
  const int grid = (int)std::min<int64_t>(mgn_num_sms(), tiles);
  const bool sv = h1_save != nullptr;
  if (pj_tma) MGN_CUDA(sv ? mgn_launch(edge_fwd_hl_kernel<true, true>, grid, NT, EDGE_HL_SMEM, stream, p)
                          : mgn_launch(edge_fwd_hl_kernel<false, true>, grid, NT, EDGE_HL_SMEM, stream, p));
  else MGN_CUDA(sv ? mgn_launch(edge_fwd_hl_kernel<true, false>, grid, NT, EDGE_HL_SMEM, stream, p)
                   : mgn_launch(edge_fwd_hl_kernel<false, false>, grid, NT, EDGE_HL_SMEM, stream, p));
  MGN_LAUNCH_CHECK();
  return mgn_launch_aggr_fixup(E, N, 32, receivers, rowptr, carry, aggr, stream);
}

// ---------------------------------------------------------------------------------------------------------------
// mgn_edge_fwd_hl for callers that cannot allocate (code under CUDA-graph capture, tight serving loops, harnesses that
// forbid cudaMalloc): the same kernels, with the carry workspace of the deterministic aggregation in a file-scope
// device array instead of a caller-provided buffer.  No counterpart in the
// reference.
//   * limit: E <= MGN_NOALLOC_MAX_EDGES (1,048,576 edges -> 32,768 slots x 512 B = 16 MiB of device memory, held by
//     every GPU that loads the library until the process exits);
//   * the array is shared by every call and keeps the last call's contents.  Nothing in it has to start at zero (a
//     carry row is only read back if this launch wrote it), so nothing is cleared; two calls that overlap on the GPU
//     (different streams, or graph branches) RACE on it -- serialise them;
//   * calls no cudaMalloc / cudaFree / allocating routine: the device address of the array is looked up once per device
//     with cudaGetSymbolAddress (the array's name is only a host-side handle in host code), and the library's lazily
//     created hang-record buffer (abi_core.cu) is used if it exists but not created from here.
static constexpr int64_t MGN_NOALLOC_MAX_EDGES = 1 << 20;
static __device__ __align__(256) float g_noalloc_carry[(MGN_NOALLOC_MAX_EDGES / 32) * H];

extern "C" int64_t mgn_noalloc_max_edges(void) { return MGN_NOALLOC_MAX_EDGES; }

extern "C" int mgn_edge_fwd_hl_noalloc(int64_t E, int64_t N, const void* hi_in, const void* lo_in, void* hi_out,
                                       void* lo_out, float* e_out, const void* P, const int32_t* senders,
                                       const int32_t* receivers, float* aggr, const void* packed, const float* b2,
                                       const float* b3, const float* b4, const float* scale, const int32_t* rowptr,
                                       int norm_dim, void* stream_) {
  if (E > MGN_NOALLOC_MAX_EDGES) {
    mgn_set_error("edge_fwd_hl_noalloc: E = %lld exceeds the built-in scratch (%lld edges); use mgn_edge_fwd_hl with a "
                  "caller-provided carry buffer", (long long)E, (long long)MGN_NOALLOC_MAX_EDGES);
    return MGN_EWORKSPACE;
  }
  constexpr int MAX_DEV = 64;
  static float* dev_ptr[MAX_DEV] = {};          // device address of g_noalloc_carry, per device
  int dev = 0;
  MGN_CUDA(cudaGetDevice(&dev));
  MGN_CHECK_ARG(dev >= 0 && dev < MAX_DEV, "edge_fwd_hl_noalloc: device ordinal out of range");
  if (dev_ptr[dev] == nullptr) {                // once per device, with that device current
    float* ptr = nullptr;
    MGN_CUDA(cudaGetSymbolAddress((void**)&ptr, g_noalloc_carry));
    MGN_CHECK_ARG(ptr != nullptr, "edge_fwd_hl_noalloc: cudaGetSymbolAddress returned null");
    dev_ptr[dev] = ptr;
  }
  mgn_set_no_alloc_scope(true);
  const int rc = mgn_edge_fwd_hl(E, N, hi_in, lo_in, hi_out, lo_out, e_out, P, senders, receivers, aggr, packed, b2, b3,
                                 b4, scale, dev_ptr[dev], rowptr, norm_dim, nullptr, nullptr, nullptr, stream_);
  mgn_set_no_alloc_scope(false);
  return rc;
}
