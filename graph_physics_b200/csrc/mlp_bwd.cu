// mlp_bwd.cu -- fused backward kernels of one MeshGraphNet message-passing step (the autograd
// of GraphNetBlock.forward, graphphysics/models/layers.py:667-746).
//
//   edge_bwd_kernel   recomputes the edge MLP from the saved step input e_k (hidden activations
//                     are never stored), RMSNorm backward, dgrad chain, weight gradients of all four
//                     layers, g_e = g_e_out + dz1 We, and the scatter of dz1 into the per-node
//                     gradients of the layer-1 partial products P (by receiver: segmented sum;
//                     by sender: red.add).
//   node_bwd_kernel   same chain for the node MLP from the fp16 hidden saves of the forward
//                     kernel; outputs g_x (direct part), g_aggr and dz1' (fp16) for node_wgrad1.
//   p_bwd_kernel      g_x += gP [W_i ; W_j], dW_i, dW_j, db1.
//   node_wgrad1_kernel dW1' = dz1'^T [x | aggr].
//
// All products run on tcgen05 (fp16 operands, fp32 accumulation in TMEM).  Weight gradients of
// three layers stay resident in TMEM for the whole kernel (3 x 128 columns) next to one 128-column
// working accumulator; the fourth (edge layer 1) is formed per tile in the working accumulator and
// summed in registers (32 per thread).  Weight tiles stream through two 32 KB shared-memory slots
// (bulk async copies from L2), the remaining five slots hold the activation / gradient tiles of the
// current 128-row tile.  512 threads per tile (mlp_v2.cuh mapping).  Gradients arrive pre-scaled by
// a power of two (processor.py) so that fp16 operands keep their precision; all of this is linear in g.
#define MGN_FILE_ID 3
#include "mlp_v2.cuh"
#include "mlp_fwd_abi.cuh"
#include <cstdlib>

using namespace mgn;

static constexpr uint32_t SLOT = 32768;

__device__ __forceinline__ void load_w_slot(uint8_t* slot, const __half* W, int widx, uint64_t* bar) {
  mbar_arrive_expect_tx(bar, SLOT);
  bulk_g2s(slot, reinterpret_cast<const uint8_t*>(W) + (size_t)widx * SLOT, SLOT, bar);
}

// RMSNorm backward for one thread's 32 columns (layers.py:49-74 differentiated):
//   y = scale * z * inv, inv = 1 / (||z|| d^-1/2 + eps);  dz = scale*g*inv - z * (sum_c scale*g*z) * inv^2 / (d r)
// acc = z - bias (TMEM), gtile row holds the upstream gradient g (fp16) and is overwritten with fp16(dz).
// Returns this warp's column sum (over its 32 rows) of g*z*inv for column c0+lane (= dL/dscale part).
__device__ __forceinline__ float rmsnorm_bwd_q(uint32_t taddr, const float4 (&bb)[8], const float4 (&sc)[8],
                                               uint8_t* gtile, float* red, const Map& m, float inv_sqrt_d,
                                               float norm_d) {
  float z[32], g[32];
  tmem_ld32(taddr, z);
  load_row_chunk_f16(gtile, 128, m.row, m.c0, g);
  tmem_ld_wait();
  float ss = 0.f, dot = 0.f;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const float b_[4] = {bb[j].x, bb[j].y, bb[j].z, bb[j].w};
    const float s_[4] = {sc[j].x, sc[j].y, sc[j].z, sc[j].w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      z[4 * j + i] += b_[i];
      ss += z[4 * j + i] * z[4 * j + i];
      dot += s_[i] * g[4 * j + i] * z[4 * j + i];
    }
  }
  ss = row_sum4(ss, red, m);
  dot = row_sum4(dot, red, m);
  const float r = sqrtf(ss) * inv_sqrt_d;
  const float inv = 1.0f / (r + RMS_EPS);
  const float k = r > 0.f ? dot * inv * inv / (norm_d * r) : 0.f;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const float s_[4] = {sc[j].x, sc[j].y, sc[j].z, sc[j].w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float zz = z[4 * j + i], gg = g[4 * j + i];
      z[4 * j + i] = s_[i] * gg * inv - zz * k;     // dL/dz
      g[4 * j + i] = gg * zz * inv;                 // dL/dscale term
    }
  }
  store_row_chunk_f16(gtile, 128, m.row, m.c0, z);
  return warp_colsum32(g, m.lane);
}

// Same with ONE exchange of the two row sums across the four column quarters: (sum z^2, sum scale g z) go as a float2
// through `scratch` ([4][128] float2 = 4 KB of shared memory that nothing else touches until the caller's next
// barrier) -- one barrier instead of four.
__device__ __forceinline__ float rmsnorm_bwd_q2(uint32_t taddr, const float4 (&bb)[8], const float4 (&sc)[8],
                                                uint8_t* gtile, float2* scratch, const Map& m, float inv_sqrt_d,
                                                float norm_d) {
  float z[32], g[32];
  tmem_ld32(taddr, z);
  load_row_chunk_f16(gtile, 128, m.row, m.c0, g);
  tmem_ld_wait();
  float ss = 0.f, dot = 0.f;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const float b_[4] = {bb[j].x, bb[j].y, bb[j].z, bb[j].w};
    const float s_[4] = {sc[j].x, sc[j].y, sc[j].z, sc[j].w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      z[4 * j + i] += b_[i];
      ss += z[4 * j + i] * z[4 * j + i];
      dot += s_[i] * g[4 * j + i] * z[4 * j + i];
    }
  }
  scratch[m.cq * 128 + m.row] = make_float2(ss, dot);
  __syncthreads();
  {
    const float2 a = scratch[m.row], b = scratch[128 + m.row], c = scratch[256 + m.row], d = scratch[384 + m.row];
    ss = (a.x + b.x) + (c.x + d.x);          // same association as row_sum4
    dot = (a.y + b.y) + (c.y + d.y);
  }
  const float r = sqrtf(ss) * inv_sqrt_d;
  const float inv = 1.0f / (r + RMS_EPS);
  const float k = r > 0.f ? dot * inv * inv / (norm_d * r) : 0.f;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const float s_[4] = {sc[j].x, sc[j].y, sc[j].z, sc[j].w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float zz = z[4 * j + i], gg = g[4 * j + i];
      z[4 * j + i] = s_[i] * gg * inv - zz * k;
      g[4 * j + i] = gg * zz * inv;
    }
  }
  store_row_chunk_f16(gtile, 128, m.row, m.c0, z);
  return warp_colsum32(g, m.lane);
}

// ===================================================================================== edge
struct EdgeBwdParams {
  long long E;
  const float* e_in;        // [E,128] step input (saved); used when e16 is null
  const __half* e16;        // fp16 tile images of e_in written by edge_fwd (32 KB per 128-row tile), or null
  const int32_t* rowmap;    // CSR position -> row of e_in / g_eout / g_e (null = identity)
  const float* g_eout;      // [E,128] or null (= 0)
  const float* g_aggr;      // [N,128]
  const __half* P;          // [N,256]
  const int32_t* senders;
  const int32_t* receivers;
  float* g_e;               // [E,128] out; may be the same buffer as g_eout
  __half* dz16_out;         // fp16 tile images of dz1 (32 KB per 128-row tile), gathered by p_bwd
  const __half* W;          // packed edge tiles We, W2, W3, W4
  const float* b2;
  const float* b3;
  const float* b4;
  const float* scale;
  float* grad;              // flat gradient block (GB_* offsets), accumulated with red.add
  long long* dbg;
  int dbg_cap;
  int dbg_skip;             // developer experiments only (MGN_DEBUG_SKIP): bit mask of memory operations to leave out
  float inv_sqrt_d, norm_d; // RMSNorm: true hidden size d and d^-1/2 (zero-padded narrower models)
};

// FAST = the processor loop's case (upstream g_e tile contiguous: no rowmap, g_eout given, e images given): the operand
// g = fp16(g_eout + g_aggr[receiver]) is built under product R2 from a bulk-loaded copy of the upstream tile (row-owner
// mapping, g_aggr chunks read directly -- rows of one receiver are consecutive), so no thread issues or waits for loads
// of the g_e stream and the "next tile's g in four pieces" schedule of the general path disappears.
template <bool FAST>
__global__ void __launch_bounds__(NT, 1) edge_bwd_kernel(const EdgeBwdParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  check_smem_alignment(smem);
  uint8_t* sWA = smem;
  uint8_t* sWB = smem + SLOT;
  uint8_t* sE = smem + 2 * SLOT;
  uint8_t* sH1 = smem + 3 * SLOT;
  uint8_t* sH2 = smem + 4 * SLOT;   // during layer 1: staging of the gathered P_i rows
  uint8_t* sH3 = smem + 5 * SLOT;   // during layer 1: staging of the gathered P_j rows
  uint8_t* sG = smem + 6 * SLOT;
  uint8_t* lin32 = sH2;             // 64 KB row-linear fp32 staging = sH2 + sH3 (dead after D2 / D3)
  Tail2* tail = reinterpret_cast<Tail2*>(smem + 7 * SLOT);
  uint64_t& bar_wA = tail->bar[0];
  uint64_t& bar_wB = tail->bar[1];
  uint64_t& bar_mma = tail->bar[2];
  uint64_t& bar_e = tail->bar[4];
  uint64_t& bar_g = tail->bar[5];
  uint64_t& bar_g2 = tail->bar[6];
  uint64_t& bar_wC = tail->bar[7];
  uint32_t& tmem_slot = tail->tmem_slot;
  const Map m = make_map();

  if (m.tid == 0) {
    mbar_init(&bar_wA, 1);
    mbar_init(&bar_wB, 1);
    mbar_init(&bar_mma, 1);
    mbar_init(&bar_e, 1);
    mbar_init(&bar_g, 1);
    mbar_init(&bar_g2, 1);
    mbar_init(&bar_wC, 1);
    mbar_init_fence();
  }
  if (m.warp == 0) tmem_alloc<512>(&tmem_slot);
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  pdl_wait();        // the previous kernel in the stream is complete: global memory may be touched from here on
  pdl_launch();      // (common.cuh) the next kernel may start its own prologue under this one
  const uint32_t tm_W = tmem_slot, tm_dW4 = tmem_slot + 128, tm_dW3 = tmem_slot + 256, tm_dW2 = tmem_slot + 384;
  const uint32_t qoff = m.lane_off + m.c0;
  const long long num_tiles = (p.E + TILE_ROWS - 1) / TILE_ROWS;
  const uint32_t uWA = smem_u32(sWA), uWB = smem_u32(sWB), uE = smem_u32(sE), uH1 = smem_u32(sH1),
                 uH2 = smem_u32(sH2), uH3 = smem_u32(sH3), uG = smem_u32(sG);
  const uint64_t pol_stream = policy_evict_first();
  const uint64_t pol_keep = policy_evict_last();
  uint32_t phase = 0;
  float dsc = 0.f;
  float db2a = 0.f, db2b = 0.f, db3a = 0.f, db3b = 0.f, db4a = 0.f, db4b = 0.f;
  if (m.tid == 0) {
    load_w_slot(sWA, p.W, 0, &bar_wA);
    load_w_slot(sWB, p.W, 1, &bar_wB);
    // FAST: the e image travels through sG (dead from D4 of one tile to the g build of the next), which frees sE as
    // a THIRD weight slot: W2 and W3 stay resident for the whole kernel and only slot A alternates We <-> W4, so a
    // tile streams 64 KB of weights instead of 192 KB
    if (FAST) load_w_slot(sE, p.W, 2, &bar_wC);
  }
  const uint32_t uW3 = FAST ? uE : uWA;           // where W3 sits for R3 / D3
  const uint32_t uW4 = FAST ? uWA : uWB;          // where W4 sits for R4 / D4
  uint8_t* const sEimg = FAST ? sG : sE;          // where the e image lands
  const uint32_t uEimg = FAST ? uG : uE;
  auto mma_wait = [&]() {
    mbar_wait(&bar_mma, phase);
    phase ^= 1;
    fence_after_sync();
  };
  auto tile_ready = [&]() {
    fence_async_smem();
    fence_before_sync();
    __syncthreads();
  };
  // Indices, requested one tile ahead.  Coalesced mapping (lane i < 8 <-> tile row 8*warp + i): dst, rcv, snd.
  // Epilogue mapping (this thread's tile row m.row): own = row of g_e (used with a rowmap only).
  struct Idx { int dst, rcv, snd, own, rcv_o; };
  auto load_idx = [&](long long tile_) -> Idx {
    Idx ix{-1, -1, -1, -1, -1};
    const long long g = tile_ * TILE_ROWS + m.warp * 8 + (m.lane & 7);
    const long long go = tile_ * TILE_ROWS + m.row;
    if (tile_ < num_tiles && g < p.E) {
      ix.dst = p.rowmap ? __ldg(p.rowmap + g) : (int)g;
      ix.rcv = __ldg(p.receivers + g);
      ix.snd = __ldg(p.senders + g);
    }
    if (p.rowmap && tile_ < num_tiles && go < p.E) ix.own = __ldg(p.rowmap + go);
    if (tile_ < num_tiles && go < p.E) ix.rcv_o = __ldg(p.receivers + go);   // epilogue row's receiver (P_i gather)
    return ix;
  };
  const int hw = m.lane >> 4, q = m.lane & 15;
  // P_j[sender] rows (a true gather) are staged through sH1, coalesced: half a warp per 256-byte row.  P_i[receiver]
  // rows are read by the epilogue threads directly: consecutive tile rows share their receiver (CSR order), so a
  // warp's request touches only a handful of distinct rows.
  auto stage_pj = [&](const Idx& ix) {       // asynchronous (LDGSTS); cp_async_wait_all() + barrier before use
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int rc = __shfl_sync(0xffffffffu, ix.rcv, 2 * i + hw);
      const int sn = __shfl_sync(0xffffffffu, ix.snd, 2 * i + hw);
      cp_async16(uH1 + tile_off(128, m.warp * 8 + 2 * i + hw, 8 * q),
                 p.P + (long long)(rc >= 0 ? sn : 0) * 256 + 128 + 8 * q, rc >= 0 ? 16u : 0u, pol_keep);
    }
  };
  // g = g_eout + g_aggr[receiver] -> sG (fp16), coalesced, in four pieces of two rows per warp.  The loads of a piece
  // are issued at the start of one phase and its shared-memory stores happen at the start of the next one, so a whole
  // product + epilogue hides their latency: pieces 0..3 of tile t+1 are issued at D3, D2, D1 and the g_e phase of tile t
  // (sG is dead after D4) and the last one lands at R1 of tile t+1 (g is first read by R4).
  struct GPiece { float4 a[2], b[2]; };
  auto g_issue = [&](const Idx& ix, int piece) -> GPiece {
    GPiece r;
    const long long r0 = __shfl_sync(0xffffffffu, ix.rcv, 2 * piece), r1 = __shfl_sync(0xffffffffu, ix.rcv, 2 * piece + 1);
    const long long d0 = __shfl_sync(0xffffffffu, ix.dst, 2 * piece), d1 = __shfl_sync(0xffffffffu, ix.dst, 2 * piece + 1);
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    r.a[0] = (p.g_eout && d0 >= 0) ? ld_f4_hint(p.g_eout + d0 * H + 4 * m.lane, pol_stream) : z;
    r.a[1] = (p.g_eout && d1 >= 0) ? ld_f4_hint(p.g_eout + d1 * H + 4 * m.lane, pol_stream) : z;
    r.b[0] = r0 >= 0 ? ld_f4_hint(p.g_aggr + r0 * H + 4 * m.lane, pol_keep) : z;
    if (r1 == r0) r.b[1] = r.b[0];      // both edges of one receiver: fetched once (warp-uniform)
    else r.b[1] = r1 >= 0 ? ld_f4_hint(p.g_aggr + r1 * H + 4 * m.lane, pol_keep) : z;
    return r;
  };
  auto g_store = [&](const GPiece& r, int piece) {
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      uint2 pk;
      pk.x = pack_h2(r.a[i].x + r.b[i].x, r.a[i].y + r.b[i].y);
      pk.y = pack_h2(r.a[i].z + r.b[i].z, r.a[i].w + r.b[i].w);
      *reinterpret_cast<uint2*>(sG + tile_off(128, m.warp * 8 + 2 * piece + i, 4 * m.lane)) = pk;
    }
  };
  // e tile: the forward kernel's fp16 tile image (one 32 KB bulk copy), else converted from the fp32 rows
  auto issue_e16 = [&](long long tile_) {      // thread 0
    mbar_arrive_expect_tx(&bar_e, SLOT);
    bulk_g2s(sEimg, reinterpret_cast<const uint8_t*>(p.e16) + (size_t)tile_ * SLOT, SLOT, &bar_e);
  };
  Timeline tl;
  tl.init(m.tid == 0 && blockIdx.x == 0, p.dbg, p.dbg_cap);

  // developer experiment: de-phase the CTAs (MGN_DEBUG_SKIP bits 16.. = delay unit in 256-cycle steps)
  if ((p.dbg_skip >> 16) != 0) {
    const long long t0 = clock64(), d = (long long)(blockIdx.x & 7) * ((p.dbg_skip >> 16) << 8);
    while (clock64() - t0 < d) {}
  }
  // ---- prologue: g and e of the first tile
  Idx cur = load_idx(blockIdx.x);
  if (p.e16 != nullptr && m.tid == 0) issue_e16(blockIdx.x);
  stage_pj(cur);
  if (!FAST) {
#pragma unroll
    for (int q4 = 0; q4 < 4; ++q4) g_store(g_issue(cur, q4), q4);
  }
  GPiece gq;                            // the piece in flight (next tile's g)
  bool gq_pending = false;              // first tile: nothing left over from a previous tile
  auto tile_rows_bytes = [&](long long tile_) -> uint32_t {
    const long long r0 = tile_ * TILE_ROWS;
    return (uint32_t)(((p.E - r0) < TILE_ROWS ? (p.E - r0) : TILE_ROWS) * 512);
  };
  Idx nxt = load_idx((long long)blockIdx.x + gridDim.x);

  uint32_t it = 0;
  for (long long tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
    tl.stamp();   // 0
    const bool has_next = tile + gridDim.x < num_tiles;
    if (p.e16 == nullptr) {
      load8_f32_to_tile(sE, 128, 0, p.e_in, m,
                        [&](int i) -> long long { return (long long)__shfl_sync(0xffffffffu, cur.dst, i); }, pol_stream);
      tile_ready();
    }
    tl.stamp();   // 1

    if (!FAST && gq_pending) g_store(gq, 3);       // last piece of this tile's g
    // ---- R1: z1 = e We^T + P_i[recv] + P_j[send]
    if (m.tid == 0) {
      if (it == 0) mbar_wait(&bar_wA, 0);
      if (p.e16 != nullptr) mbar_wait(&bar_e, it & 1);
      fence_after_sync();
      umma_tile(tm_W, uEimg, 128, 0, uWA, 128, 0, 0, 128, 8, 0);
      mma_commit(&bar_mma);
    }
    {
      // under the product: P_j rows -> sH1 (staging, overwritten in place by h1), P_i rows -> registers
      uint4 pi[4];
      const bool ok = cur.rcv_o >= 0;
      const __half* si = p.P + (long long)(ok ? cur.rcv_o : 0) * 256 + m.c0;
#pragma unroll
      for (int j = 0; j < 4; ++j) pi[j] = ok ? ld_u4_hint(si + 8 * j, pol_keep) : make_uint4(0, 0, 0, 0);
      cp_async_wait_all();
      __syncthreads();      // the staged P_j rows (issued during the previous tile) are visible to every thread
      mma_wait();
      if (m.tid == 0) {
        if (FAST) {             // upstream rows of this tile -> lin32, once the previous tile's g_e store has left it
          bulk_wait_read0();    // (not earlier: while thread 0 sits here the other warps wait at the next barrier)
          mbar_arrive_expect_tx(&bar_g2, tile_rows_bytes(tile));
          bulk_g2s(lin32, p.g_eout + tile * TILE_ROWS * H, tile_rows_bytes(tile), &bar_g2);
        }
        load_w_slot(sWA, p.W, FAST ? 3 : 2, &bar_wA);   // W3 -> A (FAST: W4 -> A, W3 is resident in slot C)
        // sE is dead: the next tile's e image (FAST: requested later, at D4 -- here it would compete with the
        // upstream tile the g build is about to wait for)
        if (!FAST && has_next && p.e16 != nullptr) issue_e16(tile + gridDim.x);
      }
      float v[32], b[32];
      tmem_ld32(tm_W + qoff, v);
      load_row_chunk_f16(sH1, 128, m.row, m.c0, b);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float a[8];
        unpack8(pi[j], a);
#pragma unroll
        for (int k = 0; k < 8; ++k) v[8 * j + k] = fmaxf(v[8 * j + k] + a[k] + b[8 * j + k], 0.f);
      }
      store_row_chunk_f16(sH1, 128, m.row, m.c0, v);
    }
    if (!FAST && m.tid == 0) bulk_wait_read0();    // the previous tile's g_e store has read lin32 (= sH2 + sH3)
    // FAST: this row's g_aggr chunks (chunk order rotated by lane & 7, see the g build under R2), requested a barrier,
    // the prefetches and a product issue ahead of their use
    float4 gag[8];
    if (FAST) {
      const int kx = m.lane & 7;
      const bool ok = cur.rcv_o >= 0;
      const float* ga = p.g_aggr + (long long)(ok ? cur.rcv_o : 0) * H + m.c0;
#pragma unroll
      for (int c = 0; c < 8; ++c) gag[c] = ok ? ld_f4_hint(ga + 4 * (c ^ kx), pol_keep) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    tile_ready();
    tl.stamp();   // 2 R1 done
    // L2 prefetch of the next tile's streamed rows: the upstream tile is contiguous without a rowmap (one bulk
    // prefetch by one thread); g_aggr rows once per distinct receiver (consecutive rows share it)
    {
      const int nd = __shfl_sync(0xffffffffu, nxt.dst, m.lane & 7);
      const int nr = __shfl_sync(0xffffffffu, nxt.rcv, m.lane & 7);
      const int np = __shfl_sync(0xffffffffu, nxt.rcv, (m.lane & 7) ? (m.lane & 7) - 1 : 0);
      const int part = m.lane >> 3;           // 4 lanes per row: one 128-byte line each
      if (nd >= 0) {
        if (p.g_eout && p.rowmap) prefetch_l2(reinterpret_cast<const char*>(p.g_eout + (long long)nd * H) + 128 * part);
        if ((m.lane & 7) == 0 || nr != np)
          prefetch_l2(reinterpret_cast<const char*>(p.g_aggr + (long long)nr * H) + 128 * part);
      }
      if (m.tid == 0 && p.g_eout && !p.rowmap && has_next) {
        const long long nrow0 = (tile + gridDim.x) * TILE_ROWS;
        const long long nrows = (p.E - nrow0) < TILE_ROWS ? (p.E - nrow0) : TILE_ROWS;
        bulk_prefetch_l2(p.g_eout + nrow0 * H, (uint32_t)(nrows * 512));
      }
    }
    // ---- R2
    {
      if (m.tid == 0) {
        if (it == 0) mbar_wait(&bar_wB, 0);
        fence_after_sync();
        umma_tile(tm_W, uH1, 128, 0, uWB, 128, 0, 0, 128, 8, 0);
        mma_commit(&bar_mma);
      }
      if (FAST) {
        // g = fp16(g_eout + g_aggr[receiver]) for this thread's 32 columns.  lin32 is row-linear with the 16-byte
        // chunks of a row rotated by (lane & 7) on access (bank-conflict free); the same rotation picks the g_aggr
        // chunk and the destination columns, so nothing is permuted in registers.
        const int kx = m.lane & 7;
        const bool ok = cur.rcv_o >= 0;
        const float4 (&a)[8] = gag;
        mbar_wait(&bar_g2, it & 1);
        const uint8_t* base = lin32 + m.row * 512 + m.c0 * 4;
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          const float4 u = ok ? *reinterpret_cast<const float4*>(base + 16 * (c ^ kx)) : make_float4(0.f, 0.f, 0.f, 0.f);
          uint2 pk;
          pk.x = pack_h2(u.x + a[c].x, u.y + a[c].y);
          pk.y = pack_h2(u.z + a[c].z, u.w + a[c].w);
          *reinterpret_cast<uint2*>(sG + tile_off(128, m.row, m.c0 + 4 * (c ^ kx))) = pk;
        }
        __syncthreads();        // every thread has read lin32 before the epilogue below writes sH2
      }
      float4 bb[8];
      load_vec32(p.b2, m.c0, bb);
      mma_wait();
      if (!FAST && m.tid == 0) load_w_slot(sWB, p.W, 3, &bar_wB);     // W4 -> B
      epi_bias_relu_q(tm_W + qoff, bb, sH2, m, nullptr, 0);
      tile_ready();
    }
    // ---- R3
    {
      if (m.tid == 0) {
        if (!FAST) mbar_wait(&bar_wA, 1);
        else if (it == 0) mbar_wait(&bar_wC, 0);
        fence_after_sync();
        umma_tile(tm_W, uH2, 128, 0, uW3, 128, 0, 0, 128, 8, 0);
        mma_commit(&bar_mma);
      }
      float4 bb[8];
      load_vec32(p.b3, m.c0, bb);
      mma_wait();
      epi_bias_relu_q(tm_W + qoff, bb, sH3, m, nullptr, 0);
      tile_ready();
    }
    tl.stamp();   // 3 R2, R3 done
    // ---- R4 + RMSNorm backward
    {
      if (m.tid == 0) {
        mbar_wait(FAST ? &bar_wA : &bar_wB, 1);
        fence_after_sync();
        umma_tile(tm_W, uH3, 128, 0, uW4, 128, 0, 0, 128, 8, 0);
        mma_commit(&bar_mma);
      }
      float4 bb[8], sc[8];
      load_vec32(p.b4, m.c0, bb);
      load_vec32(p.scale, m.c0, sc);
      mma_wait();
      dsc += rmsnorm_bwd_q(tm_W + qoff, bb, sc, sG, tail->red, m, p.inv_sqrt_d, p.norm_d);
      tile_ready();
    }
    tl.stamp();   // 4 R4 + norm done
    // ---- D4: g_h3 = dz4 W4 ; dW4 += dz4^T h3
    if (m.tid == 0) {
      fence_after_sync();
      umma_tile(tm_W, uG, 128, 0, uW4, 128, 1, 0, 128, 8, 0);
      umma_tile(tm_dW4, uG, 128, 1, uH3, 128, 1, 0, 128, 8, it > 0);
      mma_commit(&bar_mma);
    }
    tile_colsum_q(sG, m.tid, db4a, db4b);
    mma_wait();
    if (m.tid == 0) {
      if (!FAST) load_w_slot(sWB, p.W, 1, &bar_wB);     // W2 -> B
      if (FAST) {
        load_w_slot(sWA, p.W, 0, &bar_wA);                 // W4 is done: We -> A for D1 (and the next tile's R1)
      }
    }
    epi_relu_mask_q(tm_W + qoff, sH3, m);
    tile_ready();
    // ---- D3
    if (m.tid == 0) {
      // FAST: sG is dead (D4's products are done and every thread is past its column sums of it, barrier above):
      // the next tile's e image lands there
      if (FAST && has_next) issue_e16(tile + gridDim.x);
      fence_after_sync();
      umma_tile(tm_W, uH3, 128, 0, uW3, 128, 1, 0, 128, 8, 0);
      umma_tile(tm_dW3, uH3, 128, 1, uH2, 128, 1, 0, 128, 8, it > 0);
      mma_commit(&bar_mma);
    }
    if (!FAST && has_next) gq = g_issue(nxt, 0);   // sG is dead: next tile's g, piece 0
    tile_colsum_q(sH3, m.tid, db3a, db3b);
    mma_wait();
    if (!FAST && m.tid == 0) load_w_slot(sWA, p.W, 0, &bar_wA);     // We -> A
    epi_relu_mask_q(tm_W + qoff, sH2, m);
    tile_ready();
    // ---- D2
    if (m.tid == 0) {
      if (!FAST) mbar_wait(&bar_wB, 0);
      fence_after_sync();
      umma_tile(tm_W, uH2, 128, 0, uWB, 128, 1, 0, 128, 8, 0);
      umma_tile(tm_dW2, uH2, 128, 1, uH1, 128, 1, 0, 128, 8, it > 0);
      mma_commit(&bar_mma);
    }
    if (!FAST && has_next) {
      g_store(gq, 0);
      gq = g_issue(nxt, 1);
    }
    tile_colsum_q(sH2, m.tid, db2a, db2b);
    mma_wait();
    epi_relu_mask_q(tm_W + qoff, sH1, m);    // sH1 <- dz1
    tile_ready();
    tl.stamp();   // 5 D4..D2 done
    // upstream rows of this tile for the residual add, straight into the (now dead) sH2 + sH3 as a linear fp32 tile
    const long long row0 = tile * TILE_ROWS;
    const uint32_t tile_bytes = (uint32_t)(((p.E - row0) < TILE_ROWS ? (p.E - row0) : TILE_ROWS) * 512);
    // in place (g_e is the buffer g_eout lives in: the processor loop): g_e += dz1 We is ONE bulk reduce-add of the
    // accumulator tile, performed at L2 -- the upstream rows are not fetched a second time (64 KB less shared-memory
    // fill per tile, no wait for them, no read-modify-write of the staging tile).  Every element receives exactly one
    // add, so the result does not depend on any ordering.
    const bool red_inplace = FAST && p.g_e == p.g_eout;
    if (m.tid == 0 && p.g_eout != nullptr && p.rowmap == nullptr && !red_inplace) {
      mbar_arrive_expect_tx(&bar_g, tile_bytes);
      bulk_g2s(lin32, p.g_eout + row0 * H, tile_bytes, &bar_g);
    }
    // ---- the dz1 tile goes to global memory as it is (one bulk copy of the fp16 operand tile): dp_gather sums it per
    //      receiver and per sender (nothing is scattered with atomics here) and edge_wgrad1 forms dWe = dz1^T e from it
    if (m.tid == 0 && !(p.dbg_skip & 64)) {
      bulk_s2g(reinterpret_cast<uint8_t*>(p.dz16_out) + (size_t)tile * SLOT, sH1, SLOT, pol_keep);
      bulk_commit();
    }
    // ---- D1: dz1 We; under it: the next tile's g (coalesced loads)
    if (m.tid == 0) {
      mbar_wait(&bar_wA, 0);
      fence_after_sync();
      umma_tile(tm_W, uH1, 128, 0, uWA, 128, 1, 0, 128, 8, 0);
      mma_commit(&bar_mma);
    }
    const Idx done = cur;
    cur = nxt;
    if (!FAST && has_next) {
      g_store(gq, 1);
      gq = g_issue(cur, 2);
    }
    nxt = load_idx(tile + 2 * (long long)gridDim.x);
    mma_wait();
    if (!FAST && has_next) {
      g_store(gq, 2);
      gq = g_issue(cur, 3);
    }
    gq_pending = has_next;
    tl.stamp();   // 7 D1 done, next g loaded
    // ---- g_e = g_eout + dz1 We
    if (p.rowmap == nullptr) {
      // accumulator added to the upstream tile in lin32 (or written there), then one bulk store of the tile
      if (p.g_eout != nullptr && !red_inplace) mbar_wait(&bar_g, it & 1);
      float v[32];
      tmem_ld32(tm_W + qoff, v);
      tmem_ld_wait();
      if (p.g_eout != nullptr && !red_inplace) add_row32_linear(lin32, m, v);
      else store_row32_linear(lin32, m, v);
      if (m.tid == 0) bulk_wait_read0();    // the dz1 image copy has read sH1
      tile_ready();
      if (m.tid == 0 && !(p.dbg_skip & 8)) {
        if (red_inplace) bulk_red_add_f32(p.g_e + row0 * H, lin32, tile_bytes, pol_stream);
        else bulk_s2g(p.g_e + row0 * H, lin32, tile_bytes, pol_stream);
        bulk_commit();
      }
    } else {
      // caller's edge order: each thread moves its 32 columns of row `own` with 256-bit accesses
      float v[32];
      tmem_ld32(tm_W + qoff, v);
      tmem_ld_wait();
      if (done.own >= 0) {
        const long long o = (long long)done.own * H + m.c0;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          if (p.g_eout != nullptr) {
            const f8 old = ld_f8_hint(p.g_eout + o + 8 * j, pol_stream);
#pragma unroll
            for (int k = 0; k < 8; ++k) v[8 * j + k] += old.v[k];
          }
          st_f8_hint(p.g_e + o + 8 * j, &v[8 * j], pol_stream);
        }
      }
      fence_before_sync();
      if (m.tid == 0) bulk_wait_read0();    // the dz1 image copy has read sH1
      __syncthreads();      // every thread has read the accumulator before the next tile's first product
    }
    if (has_next) stage_pj(cur);            // sH1 is dead: the next tile's P_j rows, asynchronously
    tl.stamp();   // 8 g_e issued
  }
  if (m.tid == 0) bulk_wait0();
  tl.stamp();   // loop done
  // ---------------- flush the resident weight gradients, bias and scale gradients
  fence_after_sync();
  __syncthreads();          // thread 0 has seen the last g_e store leave lin32
  // (every bulk load of the loop was consumed inside it, the stores were waited for above: all slots are dead)
  flush_acc_bulk3(tm_dW4 + qoff, tm_dW3 + qoff, tm_dW2 + qoff, p.grad + GB_E_W4, p.grad + GB_E_W3, p.grad + GB_E_W2,
                  smem, m, pol_keep);
  const int cp = m.tid & 63;
  atomicAdd(p.grad + GB_E_B4 + 2 * cp, db4a);
  atomicAdd(p.grad + GB_E_B4 + 2 * cp + 1, db4b);
  atomicAdd(p.grad + GB_E_B3 + 2 * cp, db3a);
  atomicAdd(p.grad + GB_E_B3 + 2 * cp + 1, db3b);
  atomicAdd(p.grad + GB_E_B2 + 2 * cp, db2a);
  atomicAdd(p.grad + GB_E_B2 + 2 * cp + 1, db2b);
  atomicAdd(p.grad + GB_E_SC + m.c0 + m.lane, dsc);
  tl.stamp();   // flush issued

  fence_before_sync();
  __syncthreads();
  if (m.warp == 0) tmem_dealloc<512>(tmem_slot);
}

// ===================================================================================== edge, from saved activations
// The processor loop's edge backward when the forward kept the edge MLP's hidden activations (mgn_edge_fwd_hl h*_save):
// no recompute of layers 1-3, no gathers of P, no e image -- the tile starts at z4 = h3 W4^T (needed by the RMSNorm
// backward), then the same dgrad / wgrad chain as edge_bwd_kernel.  g_e is updated IN PLACE: the upstream rows are
// read once (bulk copy, for the operand g = fp16(g_e + g_aggr[receiver])) and g_e += dz1 We leaves as one bulk
// reduce-add performed at L2.  Shared memory (7 x 32 KB):
//   sA  W4 (R4, D4) <-> W2 (D2)        sB  W3 (resident)
//   sH1 h1 -> dz1 | sC We              together: landing zone of the upstream fp32 tile (64 KB) before h1 / We arrive
//   sH3 h3 -> dz3
//   sH2 h2 -> dz2 | sG g -> dz4        together: row-linear fp32 staging of the g_e tile (64 KB) after D2
// The product R4 of the NEXT tile is issued before the tile boundary (its operands h3, W4 are loaded as soon as their
// slots die), so it runs under the reduce-add's read-out, the upstream wait and the build of g.
struct EdgeBwdSavedParams {
  long long E;
  const __half* h1;         // tile images of the hidden activations (32 KB per 128-row tile)
  const __half* h2;
  const __half* h3;
  const float* g_aggr;      // [N,128]
  const int32_t* receivers;
  float* g_e;               // [E,128] in: upstream gradient, out: g_e + dz1 We (in place)
  __half* dz16_out;         // tile images of dz1
  const __half* W;          // packed edge tiles We, W2, W3, W4
  const float* b4;
  const float* scale;
  float* grad;
  float inv_sqrt_d, norm_d;
  long long* dbg;           // developer timeline (MGN_TIMELINE builds)
  int dbg_cap;
};

__global__ void __launch_bounds__(NT, 1) edge_bwd_saved_kernel(const EdgeBwdSavedParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  check_smem_alignment(smem);
  uint8_t* sA = smem;
  uint8_t* sB = smem + SLOT;
  uint8_t* sH1 = smem + 2 * SLOT;
  uint8_t* sC = smem + 3 * SLOT;
  uint8_t* sH3 = smem + 4 * SLOT;
  uint8_t* sH2 = smem + 5 * SLOT;
  uint8_t* sG = smem + 6 * SLOT;
  uint8_t* lin_in = sH1;            // sH1 + sC
  uint8_t* lin_out = sH2;           // sH2 + sG
  Tail2* tail = reinterpret_cast<Tail2*>(smem + 7 * SLOT);
  uint64_t& bar_wA = tail->bar[0];
  uint64_t& bar_wB = tail->bar[1];
  uint64_t& bar_mma = tail->bar[2];
  uint64_t& bar_wC = tail->bar[3];
  uint64_t& bar_up = tail->bar[4];
  uint64_t& bar_h3 = tail->bar[5];
  uint64_t& bar_h2 = tail->bar[6];
  uint64_t& bar_h1 = tail->bar[7];
  uint32_t& tmem_slot = tail->tmem_slot;
  const Map m = make_map();
  if (m.tid == 0) {
    for (int i = 0; i < 8; ++i) mbar_init(&tail->bar[i], 1);
    mbar_init_fence();
  }
  if (m.warp == 0) tmem_alloc<512>(&tmem_slot);
  // The first tile's streams into L2 while the previous kernel drains: the hidden images were written by the forward
  // pass and g_e by the edge backward of the previous step, five launches back -- none of them by the kernel this
  // one's griddepcontrol.wait is for, so touching them here is safe.
  if (m.tid == 64 && (long long)blockIdx.x * TILE_ROWS < p.E) {
    const long long t0 = blockIdx.x;
    prefetch_tile_rows_l2(p.g_e, t0 * TILE_ROWS, p.E, 512);
    bulk_prefetch_l2(reinterpret_cast<const uint8_t*>(p.h3) + (size_t)t0 * SLOT, SLOT);
    bulk_prefetch_l2(reinterpret_cast<const uint8_t*>(p.h2) + (size_t)t0 * SLOT, SLOT);
    bulk_prefetch_l2(reinterpret_cast<const uint8_t*>(p.h1) + (size_t)t0 * SLOT, SLOT);
  }
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  pdl_wait();
  pdl_launch();
  const uint32_t tm_W = tmem_slot, tm_dW4 = tmem_slot + 128, tm_dW3 = tmem_slot + 256, tm_dW2 = tmem_slot + 384;
  const uint32_t qoff = m.lane_off + m.c0;
  const long long num_tiles = (p.E + TILE_ROWS - 1) / TILE_ROWS;
  const uint32_t uA = smem_u32(sA), uB = smem_u32(sB), uC = smem_u32(sC), uH1 = smem_u32(sH1), uH2 = smem_u32(sH2),
                 uH3 = smem_u32(sH3), uG = smem_u32(sG);
  const uint64_t pol_stream = policy_evict_first();
  const uint64_t pol_keep = policy_evict_last();
  uint32_t phase = 0;
  float dsc = 0.f;
  float db2a = 0.f, db2b = 0.f, db3a = 0.f, db3b = 0.f, db4a = 0.f, db4b = 0.f;
  auto mma_wait = [&]() {
    mbar_wait(&bar_mma, phase);
    phase ^= 1;
    fence_after_sync();
  };
  auto tile_ready = [&]() {
    fence_async_smem();
    fence_before_sync();
    __syncthreads();
  };
  auto tile_rows_bytes = [&](long long tile_) -> uint32_t {
    const long long r0 = tile_ * TILE_ROWS;
    return (uint32_t)(((p.E - r0) < TILE_ROWS ? (p.E - r0) : TILE_ROWS) * 512);
  };
  // Bulk LOADS (non-blocking) are issued by thread 32, so that they overlap thread 0's product issue; bulk STORES and
  // the waits for their read-out stay with thread 0 (a bulk group can only be waited for by the thread that made it).
  const bool loader = m.tid == 32;
  auto load_img = [&](uint8_t* dst, const __half* img, long long tile_, uint64_t* bar) {      // one thread
    mbar_arrive_expect_tx(bar, SLOT);
    bulk_g2s(dst, reinterpret_cast<const uint8_t*>(img) + (size_t)tile_ * SLOT, SLOT, bar);
  };
  auto load_upstream = [&](long long tile_) {                                                 // thread 0
    mbar_arrive_expect_tx(&bar_up, tile_rows_bytes(tile_));
    bulk_g2s(lin_in, p.g_e + tile_ * TILE_ROWS * H, tile_rows_bytes(tile_), &bar_up);
  };
  Timeline tl;
  tl.init(false, nullptr, 0);
#ifdef MGN_FINE_TL           // developer build: five more stamps around the issue of R4 (tools/timeline.py edge_bwd_saved 16)
#define FINE_STAMP() tl.stamp()
#else
#define FINE_STAMP()
#endif
  auto issue_r4 = [&](uint32_t it_) {        // thread 0: z4 = h3 W4^T of the CTA's it_-th tile
    FINE_STAMP();
    mbar_wait(&bar_wA, 0);                   // W4: completions 0, 2, 4, ... of slot A
    FINE_STAMP();
    mbar_wait(&bar_h3, it_ & 1);
    FINE_STAMP();
    fence_after_sync();
    umma_tile(tm_W, uH3, 128, 0, uA, 128, 0, 0, 128, 8, 0);
    FINE_STAMP();
    mma_commit(&bar_mma);
    FINE_STAMP();
  };
  // this thread's row of the tile -> receiver (or -1 past the end); its g_aggr chunks, rotated by lane & 7 (g build)
  auto load_rcv = [&](long long tile_) -> int {
    const long long go = tile_ * TILE_ROWS + m.row;
    return (tile_ < num_tiles && go < p.E) ? __ldg(p.receivers + go) : -1;
  };
  float4 gag[8];
  auto load_gag = [&](int rcv) {
    const int kx = m.lane & 7;
    const float* ga = p.g_aggr + (long long)(rcv >= 0 ? rcv : 0) * H + m.c0;
#pragma unroll
    for (int c = 0; c < 8; ++c)
      gag[c] = rcv >= 0 ? ld_f4_hint(ga + 4 * (c ^ kx), pol_keep) : make_float4(0.f, 0.f, 0.f, 0.f);
  };

  // b4 and scale live in the tail's 2 KB scratch (unused here: the row sums go through sH1): the norm backward reads
  // them from shared memory instead of issuing 16 global loads per thread and tile behind the tile's other traffic
  float* svec = tail->red;                   // [0,128) b4, [128,256) scale
  if (m.tid < 128) {
    svec[m.tid] = __ldg(p.b4 + m.tid);
    svec[128 + m.tid] = __ldg(p.scale + m.tid);
  }
  __syncthreads();
  // ---- prologue
  const long long tile0 = blockIdx.x;
  if (loader) {
    load_w_slot(sA, p.W, 3, &bar_wA);      // W4
    load_img(sH3, p.h3, tile0, &bar_h3);
    load_upstream(tile0);
    load_w_slot(sB, p.W, 2, &bar_wB);      // W3
    load_img(sH2, p.h2, tile0, &bar_h2);
  }
  int rcv_cur = load_rcv(tile0);
  int rcv_nxt = load_rcv(tile0 + gridDim.x);
  if (m.tid == 0) issue_r4(0);
  tl.init(m.tid == 0 && blockIdx.x == 0, p.dbg, p.dbg_cap);

  uint32_t it = 0;
  for (long long tile = tile0; tile < num_tiles; tile += gridDim.x, ++it) {
    const bool has_next = tile + gridDim.x < num_tiles;
    const long long row0 = tile * TILE_ROWS;
    // ---- G: the previous tile's g_e staging (sH2 + sG) has been read out -> h2 may land, sG may be written
    tl.stamp();   // 0 tile top
    load_gag(rcv_cur);                       // in flight under the waits below
    if (it > 0) {
      // the previous tile's g_e staging left as two reduce-adds, the sG half (rows 64..127) first: only that one has
      // to be read out before g can be built; the sH2 half is waited for where h2 is requested (after the norm)
      if (m.tid == 0) bulk_wait_read1();
      __syncthreads();
    }
    tl.stamp();   // 1 staging read out
    {
      const int kx = m.lane & 7;
      const bool ok = rcv_cur >= 0;
      mbar_wait(&bar_up, it & 1);
      tl.stamp();   // 2 upstream landed
      const uint8_t* base = lin_in + m.row * 512 + m.c0 * 4;
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        const float4 u = ok ? *reinterpret_cast<const float4*>(base + 16 * (c ^ kx)) : make_float4(0.f, 0.f, 0.f, 0.f);
        uint2 pk;
        pk.x = pack_h2(u.x + gag[c].x, u.y + gag[c].y);
        pk.y = pack_h2(u.z + gag[c].z, u.w + gag[c].w);
        *reinterpret_cast<uint2*>(sG + tile_off(128, m.row, m.c0 + 4 * (c ^ kx))) = pk;
      }
    }
    tile_ready();                            // g is complete; the landing zone is free
    tl.stamp();   // 3 g built
    // ---- R4 (issued before the tile boundary) + RMSNorm backward: sG <- dz4
    {
      float4 bb[8], sc[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        bb[j] = *reinterpret_cast<const float4*>(svec + m.c0 + 4 * j);
        sc[j] = *reinterpret_cast<const float4*>(svec + 128 + m.c0 + 4 * j);
      }
      mma_wait();
      tl.stamp();   // 4 R4 product done
      // the landing zone (sH1 + sC) is idle until h1 / We are requested below: its first 4 KB carry the row sums
      dsc += rmsnorm_bwd_q2(tm_W + qoff, bb, sc, sG, reinterpret_cast<float2*>(sH1), m, p.inv_sqrt_d, p.norm_d);
      if (m.tid == 0 && it > 0) bulk_wait_read0();        // the sH2 half of the previous staging has been read out too
      tile_ready();
    }
    tl.stamp();   // 5 norm backward done
    if (loader) {                            // needed by D3 / D2 / D1, one to three products from here
      if (it > 0) load_img(sH2, p.h2, tile, &bar_h2);     // (thread 0 saw the sH2 half read out before the barrier above)
      load_img(sH1, p.h1, tile, &bar_h1);
      load_w_slot(sC, p.W, 0, &bar_wC);      // We
      if (has_next) {                        // next tile's streams into L2
        const long long nt = tile + gridDim.x;
        bulk_prefetch_l2(p.g_e + nt * TILE_ROWS * H, tile_rows_bytes(nt));
        bulk_prefetch_l2(reinterpret_cast<const uint8_t*>(p.h3) + (size_t)nt * SLOT, SLOT);
        bulk_prefetch_l2(reinterpret_cast<const uint8_t*>(p.h2) + (size_t)nt * SLOT, SLOT);
        bulk_prefetch_l2(reinterpret_cast<const uint8_t*>(p.h1) + (size_t)nt * SLOT, SLOT);
      }
    }
    // ---- D4: g_h3 = dz4 W4 ; dW4 += dz4^T h3
    if (m.tid == 0) {
      fence_after_sync();
      umma_tile(tm_W, uG, 128, 0, uA, 128, 1, 0, 128, 8, 0);
      umma_tile(tm_dW4, uG, 128, 1, uH3, 128, 1, 0, 128, 8, it > 0);
      mma_commit(&bar_mma);
    }
    tile_colsum_q(sG, m.tid, db4a, db4b);
    mma_wait();
    if (loader) load_w_slot(sA, p.W, 1, &bar_wA);            // W4 is done: W2 -> A
    epi_relu_mask_q(tm_W + qoff, sH3, m);                    // sH3 <- dz3
    tile_ready();
    tl.stamp();   // 6 D4 done
    // ---- D3
    if (m.tid == 0) {
      if (it == 0) mbar_wait(&bar_wB, 0);
      mbar_wait(&bar_h2, it & 1);
      fence_after_sync();
      umma_tile(tm_W, uH3, 128, 0, uB, 128, 1, 0, 128, 8, 0);
      umma_tile(tm_dW3, uH3, 128, 1, uH2, 128, 1, 0, 128, 8, it > 0);
      mma_commit(&bar_mma);
    }
    tile_colsum_q(sH3, m.tid, db3a, db3b);
    mma_wait();
    epi_relu_mask_q(tm_W + qoff, sH2, m);                    // sH2 <- dz2
    tile_ready();
    tl.stamp();   // 7 D3 done
    // sH3 is dead (D3 complete, every thread past its column sums): the next tile's h3
    if (loader && has_next) load_img(sH3, p.h3, tile + gridDim.x, &bar_h3);
    // ---- D2
    if (m.tid == 0) {
      mbar_wait(&bar_wA, 1);                 // W2: completions 1, 3, 5, ... of slot A
      mbar_wait(&bar_h1, it & 1);
      fence_after_sync();
      umma_tile(tm_W, uH2, 128, 0, uA, 128, 1, 0, 128, 8, 0);
      umma_tile(tm_dW2, uH2, 128, 1, uH1, 128, 1, 0, 128, 8, it > 0);
      mma_commit(&bar_mma);
    }
    tile_colsum_q(sH2, m.tid, db2a, db2b);
    mma_wait();
    if (loader && has_next) load_w_slot(sA, p.W, 3, &bar_wA);       // W2 is done: the next tile's W4
    epi_relu_mask_q(tm_W + qoff, sH1, m);                    // sH1 <- dz1
    tile_ready();
    tl.stamp();   // 8 D2 done
    if (m.tid == 0) {                        // dz1 leaves as its operand tile (dp_gather, edge_wgrad1 read the images)
      bulk_s2g(reinterpret_cast<uint8_t*>(p.dz16_out) + (size_t)tile * SLOT, sH1, SLOT, pol_keep);
      bulk_commit();
    }
    // ---- D1: dz1 We
    if (m.tid == 0) {
      mbar_wait(&bar_wC, it & 1);
      fence_after_sync();
      umma_tile(tm_W, uH1, 128, 0, uC, 128, 1, 0, 128, 8, 0);
      mma_commit(&bar_mma);
    }
    rcv_cur = rcv_nxt;
    rcv_nxt = load_rcv(tile + 2 * (long long)gridDim.x);
    mma_wait();
    tl.stamp();   // 9 D1 done
    if (m.tid == 0) {
      bulk_wait_read0();                     // the dz1 image copy has read sH1, D1 has read sH1 / sC: the landing zone
      if (has_next) load_upstream(tile + gridDim.x);      // is free -- the next tile's upstream rows, a phase early
    }
    // ---- g_e += dz1 We: accumulator -> row-linear staging (sH2 + sG, both dead) -> one bulk reduce-add at L2
    {
      float v[32];
      tmem_ld32(tm_W + qoff, v);
      tmem_ld_wait();
      store_row32_linear(lin_out, m, v);
    }
    tile_ready();
    tl.stamp();   // 10 g_e staged
    if (m.tid == 0) {
      const uint32_t nb = tile_rows_bytes(tile);
      if (nb > 64 * 512) {                   // rows 64.. (staged in sG) first, as their own group
        bulk_red_add_f32(p.g_e + (row0 + 64) * H, sG, nb - 64 * 512, pol_stream);
        bulk_commit();
      }
      bulk_red_add_f32(p.g_e + row0 * H, lin_out, nb > 64 * 512 ? 64 * 512 : nb, pol_stream);
      bulk_commit();
      if (has_next) issue_r4(it + 1);        // every thread has read the accumulator (barrier above)
    }
  }
  tl.stamp();     // loop done
  if (m.tid == 0) bulk_wait0();
  // ---------------- flush the resident weight gradients, bias and scale gradients
  fence_after_sync();
  __syncthreads();
  flush_acc_bulk3(tm_dW4 + qoff, tm_dW3 + qoff, tm_dW2 + qoff, p.grad + GB_E_W4, p.grad + GB_E_W3, p.grad + GB_E_W2,
                  smem, m, pol_keep);
  const int cp = m.tid & 63;
  atomicAdd(p.grad + GB_E_B4 + 2 * cp, db4a);
  atomicAdd(p.grad + GB_E_B4 + 2 * cp + 1, db4b);
  atomicAdd(p.grad + GB_E_B3 + 2 * cp, db3a);
  atomicAdd(p.grad + GB_E_B3 + 2 * cp + 1, db3b);
  atomicAdd(p.grad + GB_E_B2 + 2 * cp, db2a);
  atomicAdd(p.grad + GB_E_B2 + 2 * cp + 1, db2b);
  atomicAdd(p.grad + GB_E_SC + m.c0 + m.lane, dsc);
  tl.stamp();     // flush issued
  fence_before_sync();
  __syncthreads();
  if (m.warp == 0) tmem_dealloc<512>(tmem_slot);
}

// ===================================================================================== node
struct NodeBwdParams {
  long long N;
  const float* g_xout;      // [N,128]
  const __half* h1;         // fp16 hidden saves of node_fwd: tile images (32 KB per 128 rows), bulk-loaded
  const __half* h2;
  const __half* h3;
  float* g_x;               // [N,128] out: g_xout + dz1' W1'[:, :128]
  float* g_aggr;            // [N,128] out: dz1' W1'[:, 128:]
  __half* dz1;              // [N,128] out (fp16) for node_wgrad1
  const __half* W;          // packed node tiles: W1' (2 slots), W2', W3', W4'
  const float* b4;
  const float* scale;
  float* grad;
  int no_norm;              // the block has no RMSNorm (decoder): dz4 = g, no scale gradient
  float inv_sqrt_d, norm_d; // RMSNorm: true hidden size d and d^-1/2
  long long* dbg;           // developer timeline (MGN_TIMELINE builds)
  int dbg_cap;
  int32_t* ready_out;       // or null: ready_out[tile] = 1 once the tile's dz1, g_x, g_aggr rows are written (chaining)
  int32_t* ready_in;        // or null: ready_in[tile] is set by the p_bwd launch before this one when g_xout's tile is
                            // complete (both kernels give tile t to CTA t mod grid, so the static assignment stays)
};

__global__ void __launch_bounds__(NT, 1) node_bwd_kernel(const NodeBwdParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  check_smem_alignment(smem);
  uint8_t* sW[2] = {smem, smem + SLOT};
  uint8_t* sH1 = smem + 2 * SLOT;
  uint8_t* sH2 = smem + 3 * SLOT;
  uint8_t* sH3 = smem + 4 * SLOT;
  uint8_t* sG = smem + 5 * SLOT;
  uint8_t* stage = sH3;   // sH3 + sG
  Tail2* tail = reinterpret_cast<Tail2*>(smem + 6 * SLOT);
  uint64_t* bar_w = &tail->bar[0];
  uint64_t& bar_mma = tail->bar[2];
  uint64_t& bar_h = tail->bar[3];
  uint32_t& tmem_slot = tail->tmem_slot;
  const Map m = make_map();
  if (m.tid == 0) {
    mbar_init(&bar_w[0], 1);
    mbar_init(&bar_w[1], 1);
    mbar_init(&bar_mma, 1);
    mbar_init(&bar_h, 1);
    mbar_init_fence();
  }
  if (m.warp == 0) tmem_alloc<512>(&tmem_slot);
  // the first tile's hidden images (written by the forward pass, not by the kernel griddepcontrol.wait is for) into L2
  // while the previous kernel drains
  if (m.tid == 64 && (long long)blockIdx.x * TILE_ROWS < p.N) {
    bulk_prefetch_l2(reinterpret_cast<const uint8_t*>(p.h3) + (size_t)blockIdx.x * SLOT, SLOT);
    bulk_prefetch_l2(reinterpret_cast<const uint8_t*>(p.h2) + (size_t)blockIdx.x * SLOT, SLOT);
    bulk_prefetch_l2(reinterpret_cast<const uint8_t*>(p.h1) + (size_t)blockIdx.x * SLOT, SLOT);
  }
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const bool chained_in = p.ready_in != nullptr;
  if (!chained_in) pdl_wait();  // the previous kernel in the stream is complete: global memory may be touched
  pdl_launch();                 // (common.cuh) the next kernel may start its own prologue under this one
  const uint32_t tm_W = tmem_slot, tm_dW4 = tmem_slot + 128, tm_dW3 = tmem_slot + 256, tm_dW2 = tmem_slot + 384;
  const uint32_t qoff = m.lane_off + m.c0;
  const uint32_t uW[2] = {smem_u32(sW[0]), smem_u32(sW[1])};
  const uint32_t uH1 = smem_u32(sH1), uH2 = smem_u32(sH2), uH3 = smem_u32(sH3), uG = smem_u32(sG);
  const long long num_tiles = (p.N + TILE_ROWS - 1) / TILE_ROWS;
  const uint32_t my_tiles = (uint32_t)((num_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x);
  const uint32_t total_loads = 5 * my_tiles;
  const uint64_t pol_keep = policy_evict_last();
  // weight stream: load q carries tile-local weight (q % 5): 0 W4', 1 W3', 2 W2', 3 W1'[:, :128], 4 W1'[:, 128:]
  // packed order is W1'a, W1'b, W2', W3', W4' (slots 0..4)
  auto issue_load = [&](uint32_t qq) {
    if (qq < total_loads) {
      const int j = qq % 5;
      const int widx = j == 0 ? 4 : (j == 1 ? 3 : (j == 2 ? 2 : (j == 3 ? 0 : 1)));
      load_w_slot(sW[qq & 1], p.W, widx, &bar_w[qq & 1]);
    }
  };
  auto wait_load = [&](uint32_t qq) { mbar_wait(&bar_w[qq & 1], (qq >> 1) & 1); };
  uint32_t phase = 0;
  auto mma_wait = [&]() {
    mbar_wait(&bar_mma, phase);
    phase ^= 1;
    fence_after_sync();
  };
  auto tile_ready = [&]() {
    fence_async_smem();
    fence_before_sync();
    __syncthreads();
  };
  if (m.tid == 0) {
    issue_load(0);
    issue_load(1);
  }
  float dsc = 0.f;
  float db1a = 0.f, db1b = 0.f, db2a = 0.f, db2b = 0.f, db3a = 0.f, db3b = 0.f, db4a = 0.f, db4b = 0.f;
  uint32_t it = 0;
  Timeline tl;
  tl.init(m.tid == 0 && blockIdx.x == 0, p.dbg, p.dbg_cap);
  tl.stamp();     // kernel body start
  cta_trace(p.dbg, p.dbg_cap, 0, 0);
  for (long long tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
    const long long row0 = tile * TILE_ROWS;
    const uint32_t q0 = 5 * it;
    tl.stamp();   // 0 tile top
    auto rf = [&](int i) -> long long { const long long g = row0 + m.warp * 8 + i; return g < p.N ? g : -1; };
    if (m.tid == 0) {          // the three saved activation tiles arrive as they were stored: one bulk copy each
      mbar_arrive_expect_tx(&bar_h, 3 * SLOT);
      bulk_g2s(sH3, reinterpret_cast<const uint8_t*>(p.h3) + (size_t)tile * SLOT, SLOT, &bar_h);
      bulk_g2s(sH2, reinterpret_cast<const uint8_t*>(p.h2) + (size_t)tile * SLOT, SLOT, &bar_h);
      bulk_g2s(sH1, reinterpret_cast<const uint8_t*>(p.h1) + (size_t)tile * SLOT, SLOT, &bar_h);
    }
    if (chained_in) {          // the upstream rows of this tile come from the p_bwd launch that may still be running
      if (m.tid == 0) tile_acquire(p.ready_in, tile);
      __syncthreads();
      load8_f32_to_tile_cg(sG, 128, 0, p.g_xout, m, rf);
    } else {
      load8_f32_to_tile(sG, 128, 0, p.g_xout, m, rf, pol_keep);
    }
    tl.stamp();   // 1 g tile built
    mbar_wait(&bar_h, it & 1);
    tile_ready();
    tl.stamp();   // 2 h1..h3 landed
    // the next tile's inputs into L2 while this tile computes (after this tile's own loads: every CTA is at the same
    // point of its tile, so a prefetch issued with them would only double the burst)
    if (m.tid == 32 && tile + gridDim.x < num_tiles) {
      const long long nt = tile + gridDim.x;
      if (!chained_in) prefetch_tile_rows_l2(p.g_xout, nt * TILE_ROWS, p.N, 512);
      bulk_prefetch_l2(reinterpret_cast<const uint8_t*>(p.h3) + (size_t)nt * SLOT, SLOT);
      bulk_prefetch_l2(reinterpret_cast<const uint8_t*>(p.h2) + (size_t)nt * SLOT, SLOT);
      bulk_prefetch_l2(reinterpret_cast<const uint8_t*>(p.h1) + (size_t)nt * SLOT, SLOT);
    }
    // ---- R4: z4 = h3 W4'^T, RMSNorm backward
    {
      if (m.tid == 0) {
        wait_load(q0);
        fence_after_sync();
        umma_tile(tm_W, uH3, 128, 0, uW[q0 & 1], 128, 0, 0, 128, 8, 0);
        mma_commit(&bar_mma);
      }
      float4 bb[8], sc[8];
      load_vec32(p.b4, m.c0, bb);
      load_vec32(p.scale, m.c0, sc);
      mma_wait();
      if (!p.no_norm) dsc += rmsnorm_bwd_q(tm_W + qoff, bb, sc, sG, tail->red, m, p.inv_sqrt_d, p.norm_d);      // else sG already holds dz4 = g
      tile_ready();
    }
    tl.stamp();   // 3 R4 + norm
    // ---- D4
    if (m.tid == 0) {
      fence_after_sync();
      umma_tile(tm_W, uG, 128, 0, uW[q0 & 1], 128, 1, 0, 128, 8, 0);
      umma_tile(tm_dW4, uG, 128, 1, uH3, 128, 1, 0, 128, 8, it > 0);
      mma_commit(&bar_mma);
    }
    tile_colsum_q(sG, m.tid, db4a, db4b);
    mma_wait();
    if (m.tid == 0) issue_load(q0 + 2);
    epi_relu_mask_q(tm_W + qoff, sH3, m);
    tile_ready();
    tl.stamp();   // 4 D4
    // ---- D3
    if (m.tid == 0) {
      wait_load(q0 + 1);
      fence_after_sync();
      umma_tile(tm_W, uH3, 128, 0, uW[(q0 + 1) & 1], 128, 1, 0, 128, 8, 0);
      umma_tile(tm_dW3, uH3, 128, 1, uH2, 128, 1, 0, 128, 8, it > 0);
      mma_commit(&bar_mma);
    }
    tile_colsum_q(sH3, m.tid, db3a, db3b);
    mma_wait();
    if (m.tid == 0) issue_load(q0 + 3);
    epi_relu_mask_q(tm_W + qoff, sH2, m);
    tile_ready();
    tl.stamp();   // 5 D3
    // ---- D2
    if (m.tid == 0) {
      wait_load(q0 + 2);
      fence_after_sync();
      umma_tile(tm_W, uH2, 128, 0, uW[(q0 + 2) & 1], 128, 1, 0, 128, 8, 0);
      umma_tile(tm_dW2, uH2, 128, 1, uH1, 128, 1, 0, 128, 8, it > 0);
      mma_commit(&bar_mma);
    }
    tile_colsum_q(sH2, m.tid, db2a, db2b);
    mma_wait();
    if (m.tid == 0) issue_load(q0 + 4);
    epi_relu_mask_q(tm_W + qoff, sH1, m);     // sH1 <- dz1'
    tile_ready();
    tl.stamp();   // 6 D2
    // ---- D1a: g_x = g_xout + dz1' W1'[:, :128]
    if (m.tid == 0) {
      wait_load(q0 + 3);
      fence_after_sync();
      umma_tile(tm_W, uH1, 128, 0, uW[(q0 + 3) & 1], 128, 1, 0, 128, 8, 0);
      mma_commit(&bar_mma);
    }
    tile_colsum_q(sH1, m.tid, db1a, db1b);
    float4 old[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const long long g = rf(i);
      old[i] = g < 0 ? make_float4(0.f, 0.f, 0.f, 0.f)
                     : (chained_in ? __ldcg(reinterpret_cast<const float4*>(p.g_xout + g * H + 4 * m.lane))
                                   : ld_f4_hint(p.g_xout + g * H + 4 * m.lane, pol_keep));
    }
    {  // dz1' tile -> global fp16 (half-warp per row)
      const int hw = m.lane >> 4, q = m.lane & 15;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int r = m.warp * 8 + 2 * i + hw;
        const long long g = row0 + r;
        if (g < p.N)
          *reinterpret_cast<uint4*>(p.dz1 + g * H + 8 * q) =
              *reinterpret_cast<const uint4*>(sH1 + tile_off(128, r, 8 * q));
      }
    }
    mma_wait();
    if (m.tid == 0) issue_load(q0 + 5);
    acc_to_stage_q(tm_W + qoff, stage, m);
    fence_before_sync();
    __syncthreads();
    tl.stamp();   // 7 D1a staged
    // ---- D1b issued now (working accumulator free); g_x written meanwhile
    if (m.tid == 0) {
      wait_load(q0 + 4);
      fence_after_sync();
      umma_tile(tm_W, uH1, 128, 0, uW[(q0 + 4) & 1], 128, 1, 0, 128, 8, 0);
      mma_commit(&bar_mma);
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const long long g = rf(i);
      const float4 v = *reinterpret_cast<const float4*>(stage + stage_off(m.warp * 8 + i, 4 * m.lane));
      if (g >= 0)
        st_f4_hint(p.g_x + g * H + 4 * m.lane,
                   make_float4(old[i].x + v.x, old[i].y + v.y, old[i].z + v.z, old[i].w + v.w), pol_keep);
    }
    tl.stamp();   // 8 g_x written
    mma_wait();
    if (m.tid == 0) issue_load(q0 + 6);
    __syncthreads();    // all reads of the g_x staging are done
    acc_to_stage_q(tm_W + qoff, stage, m);
    fence_before_sync();
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const long long g = rf(i);
      const float4 v = *reinterpret_cast<const float4*>(stage + stage_off(m.warp * 8 + i, 4 * m.lane));
      if (g >= 0) st_f4_hint(p.g_aggr + g * H + 4 * m.lane, v, pol_keep);
    }
    __syncthreads();
    if (p.ready_out != nullptr && m.tid == 0) tile_publish(p.ready_out, tile);
    tl.stamp();   // 9 g_aggr written
  }
  tl.stamp();     // loop done
  cta_trace(p.dbg, p.dbg_cap, 0, 2);
  if (chained_in) pdl_wait();   // p_bwd is complete (its gradient flush included): what this launch's successors rely on
  fence_after_sync();
  flush_acc_bulk3(tm_dW4 + qoff, tm_dW3 + qoff, tm_dW2 + qoff, p.grad + GB_N_W4, p.grad + GB_N_W3, p.grad + GB_N_W2,
                  smem, m, pol_keep);
  const int cp = m.tid & 63;
  atomicAdd(p.grad + GB_N_B4 + 2 * cp, db4a);
  atomicAdd(p.grad + GB_N_B4 + 2 * cp + 1, db4b);
  atomicAdd(p.grad + GB_N_B3 + 2 * cp, db3a);
  atomicAdd(p.grad + GB_N_B3 + 2 * cp + 1, db3b);
  atomicAdd(p.grad + GB_N_B2 + 2 * cp, db2a);
  atomicAdd(p.grad + GB_N_B2 + 2 * cp + 1, db2b);
  atomicAdd(p.grad + GB_N_B1 + 2 * cp, db1a);
  atomicAdd(p.grad + GB_N_B1 + 2 * cp + 1, db1b);
  atomicAdd(p.grad + GB_N_SC + m.c0 + m.lane, dsc);
  tl.stamp();     // flush issued
  fence_before_sync();
  __syncthreads();
  cta_trace(p.dbg, p.dbg_cap, 0, 3);
  if (m.warp == 0) tmem_dealloc<512>(tmem_slot);
}

// ===================================================================================== P backward
// dP [N,256] fp16 = [dP_i | dP_j] is GATHERED (dp_gather_kernel) from the fp16 tile images of dz1 that edge_bwd
// wrote -- no atomics, fixed summation order: dP_i[n] = sum of dz1 over the edges n receives (contiguous CSR
// positions), dP_j[n] = sum over the edges n sends (positions listed by the sender-sorted index).  Then
//   g_x += dP [W_i ; W_j] ; dW_i = dP_i^T x ; dW_j = dP_j^T x ; db1 = colsum dP_i          (p_bwd_kernel)
struct DpGatherParams {
  long long N;
  const __half* dz16;       // tile images of dz1: tile t = 32 KB swizzled 128 x 128 fp16 (CSR positions 128t..)
  const int32_t* rowptr;    // [N+1] receiver CSR
  const int32_t* rowptr_s;  // [N+1] sender-sorted lists
  const int32_t* out_pos;   // [E] CSR positions sorted by sender
  __half* dP;               // [N,256]
};

// 16-byte chunk `l` (columns 8l..8l+7) of row `pos` of the dz1 images
__device__ __forceinline__ uint4 ld_dz16_chunk(const __half* img, int pos, int l) {
  const uint32_t r = (uint32_t)pos & 127u;
  const size_t off = (size_t)(pos >> 7) * 32768 + (size_t)(l >> 3) * 16384 + r * 128 + (((uint32_t)l & 7u) ^ (r & 7u)) * 16;
  return ldg_u4(reinterpret_cast<const uint8_t*>(img) + off);
}

// One half-warp per (node, side): a plain high-occupancy streaming kernel (no shared memory), fp32 sums in list order.
// The kernel is a chain of dependent loads (row pointers -> positions -> rows): up to 8 rows of a node are in flight
// at once (a mesh node has ~6 edges: one round for most nodes) and the next task's row pointers are requested before
// the current task's rows are summed.
__global__ void __launch_bounds__(256) dp_gather_kernel(const DpGatherParams p) {
  pdl_wait();
  pdl_launch();
  const int l = threadIdx.x & 15;
  const long long ntask = 2 * p.N;
  const long long stride = ((long long)gridDim.x * blockDim.x) >> 4;
  long long task = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 4;
  int beg = 0, end = 0;
  if (task < ntask) {
    const int32_t* rp = (task & 1) ? p.rowptr_s : p.rowptr;
    beg = __ldg(rp + (task >> 1));
    end = __ldg(rp + (task >> 1) + 1);
  }
  for (; task < ntask; task += stride) {
    const long long n = task >> 1;
    const int side = (int)(task & 1);
    const long long nt = task + stride;
    int nbeg = 0, nend = 0;
    if (nt < ntask) {
      const int32_t* rp = (nt & 1) ? p.rowptr_s : p.rowptr;
      nbeg = __ldg(rp + (nt >> 1));
      nend = __ldg(rp + (nt >> 1) + 1);
    }
    float acc[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[k] = 0.f;
#pragma unroll 1
    for (int j = beg; j < end; j += 8) {
      int pos[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) pos[u] = (j + u < end) ? (side ? __ldg(p.out_pos + j + u) : j + u) : -1;
      uint4 v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) v[u] = pos[u] >= 0 ? ld_dz16_chunk(p.dz16, pos[u], l) : make_uint4(0, 0, 0, 0);
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        float f[8];
        unpack8(v[u], f);
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[k] += f[k];
      }
    }
    uint4 pk;
    pk.x = pack_h2(acc[0], acc[1]); pk.y = pack_h2(acc[2], acc[3]);
    pk.z = pack_h2(acc[4], acc[5]); pk.w = pack_h2(acc[6], acc[7]);
    *reinterpret_cast<uint4*>(p.dP + n * 256 + side * 128 + 8 * l) = pk;
    beg = nbeg;
    end = nend;
  }
}

struct PBwdParams {
  long long N;
  const __half* dP;         // [N,256] fp16
  const float* x;
  float* g_x;               // read-modify-write
  const __half* Wcat;       // [256 x 128] panel tile (64 KB)
  float* grad;
  int32_t* ready_out;       // or null: ready_out[tile] = 1 once the tile's g_x rows are written (chaining with node_bwd)
};

__global__ void __launch_bounds__(NT, 1) p_bwd_kernel(const PBwdParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  check_smem_alignment(smem);
  uint8_t* sW = smem;                 // 64 KB
  uint8_t* sGP = smem + 2 * SLOT;     // 128 x 256 fp16 = 64 KB ; later the fp32 staging of g_x
  uint8_t* sX = smem + 4 * SLOT;      // 32 KB
  Tail2* tail = reinterpret_cast<Tail2*>(smem + 5 * SLOT);
  uint64_t& wbar = tail->bar[0];
  uint64_t& bar_mma = tail->bar[1];
  uint32_t& tmem_slot = tail->tmem_slot;
  const Map m = make_map();
  if (m.tid == 0) {
    mbar_init(&wbar, 1);
    mbar_init(&bar_mma, 1);
    mbar_init_fence();
  }
  if (m.warp == 0) tmem_alloc<512>(&tmem_slot);
  // x (forward pass) and g_x (node_bwd, four launches back) of the first tile into L2 while the previous kernel drains
  if (m.tid == 64) {
    prefetch_tile_rows_l2(p.x, (long long)blockIdx.x * TILE_ROWS, p.N, 512);
    prefetch_tile_rows_l2(p.g_x, (long long)blockIdx.x * TILE_ROWS, p.N, 512);
  }
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  pdl_wait();        // the previous kernel in the stream is complete: global memory may be touched from here on
  pdl_launch();      // (common.cuh) the next kernel may start its own prologue under this one
  const uint32_t qoff = m.lane_off + m.c0;
  const uint32_t tm_W = tmem_slot, tm_dWi = tmem_slot + 128, tm_dWj = tmem_slot + 256;
  if (m.tid == 0) {
    mbar_arrive_expect_tx(&wbar, 2 * SLOT);
    bulk_g2s(sW, p.Wcat, SLOT, &wbar);
    bulk_g2s(sW + SLOT, reinterpret_cast<const uint8_t*>(p.Wcat) + SLOT, SLOT, &wbar);
  }
  const uint32_t uW = smem_u32(sW), uGP = smem_u32(sGP), uX = smem_u32(sX);
  const long long num_tiles = (p.N + TILE_ROWS - 1) / TILE_ROWS;
  const uint64_t pol_once = policy_evict_first(), pol_keep = policy_evict_last();
  uint32_t phase = 0, it = 0;
  float db1a = 0.f, db1b = 0.f;
  for (long long tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
    const long long row0 = tile * TILE_ROWS;
    auto rf = [&](int i) -> long long { const long long g = row0 + m.warp * 8 + i; return g < p.N ? g : -1; };
    // ---- dP rows (512 bytes of fp16) -> operand tile: half a warp per 256-byte half row
    {
      const int hw = m.lane >> 4, q = m.lane & 15;
      uint4 v[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const long long g = rf(i);
        v[i] = g >= 0 ? ld_u4_hint(p.dP + g * 256 + hw * 128 + 8 * q, pol_once) : make_uint4(0, 0, 0, 0);
      }
#pragma unroll
      for (int i = 0; i < 8; ++i)
        *reinterpret_cast<uint4*>(sGP + tile_off(128, m.warp * 8 + i, hw * 128 + 8 * q)) = v[i];
    }
    load8_f32_to_tile(sX, 128, 0, p.x, m, rf, pol_keep);
    fence_async_smem();
    fence_before_sync();
    __syncthreads();
    if (m.tid == 0) {
      if (it == 0) mbar_wait(&wbar, 0);
      fence_after_sync();
      // g_x part: [128 x 256] x [256 x 128]; B = MN-major view of the 256-row weight tile
      umma_tile(tm_W, uGP, 128, 0, uW, 256, 1, 0, 128, 16, 0);
      // dW_i, dW_j: A = MN-major views of the two column halves of the gP tile
      const uint32_t idesc = make_idesc(128, 128, 1, 1);
      for (uint32_t ks = 0; ks < 8; ++ks)
        mma_f16(tm_dWi, desc_mnmajor(uGP, 128, ks, 0), desc_mnmajor(uX, 128, ks, 0), idesc, (ks > 0 || it > 0) ? 1u : 0u);
      for (uint32_t ks = 0; ks < 8; ++ks)
        mma_f16(tm_dWj, desc_mnmajor(uGP, 128, ks, 128), desc_mnmajor(uX, 128, ks, 0), idesc,
                (ks > 0 || it > 0) ? 1u : 0u);
      mma_commit(&bar_mma);
    }
    tile_colsum_q(sGP, m.tid, db1a, db1b);       // columns 0..127 = dP_i
    float4 old[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const long long g = rf(i);
      old[i] = g >= 0 ? ld_f4_hint(p.g_x + g * H + 4 * m.lane, pol_keep) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    if (m.tid == 32 && tile + gridDim.x < num_tiles) {      // the next tile's inputs into L2, after this tile's own loads
      const long long n0 = (tile + gridDim.x) * TILE_ROWS;
      prefetch_tile_rows_l2(p.dP, n0, p.N, 512);
      prefetch_tile_rows_l2(p.x, n0, p.N, 512);
      prefetch_tile_rows_l2(p.g_x, n0, p.N, 512);
    }
    mbar_wait(&bar_mma, phase);
    phase ^= 1;
    fence_after_sync();
    __syncthreads();      // every thread's column sums over the fp16 gP tile are done: it becomes the staging
    acc_to_stage_q(tm_W + qoff, sGP, m);
    fence_before_sync();
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const long long g = rf(i);
      const float4 v = *reinterpret_cast<const float4*>(sGP + stage_off(m.warp * 8 + i, 4 * m.lane));
      if (g >= 0)
        st_f4_hint(p.g_x + g * H + 4 * m.lane,
                   make_float4(old[i].x + v.x, old[i].y + v.y, old[i].z + v.z, old[i].w + v.w), pol_keep);
    }
    __syncthreads();
    if (p.ready_out != nullptr && m.tid == 0) tile_publish(p.ready_out, tile);
  }
  fence_after_sync();
  // dW_i -> edge_block.0.weight[:, 128:256], dW_j -> [:, 256:384]
  flush_acc_bulk(tm_dWi + qoff, p.grad + GB_E_W1 + 128, 384, sGP, m, pol_keep);
  flush_acc_bulk(tm_dWj + qoff, p.grad + GB_E_W1 + 256, 384, sGP, m, pol_keep);
  const int cp = m.tid & 63;
  atomicAdd(p.grad + GB_E_B1 + 2 * cp, db1a);
  atomicAdd(p.grad + GB_E_B1 + 2 * cp + 1, db1b);
  fence_before_sync();
  __syncthreads();
  if (m.warp == 0) tmem_dealloc<512>(tmem_slot);
}

// ===================================================================================== node layer-1 wgrad
// dW1' [128, 256] += dz1'^T [x | aggr].  Two 96 KB operand sets: the loads of tile i+1 run under the MMAs of tile i.
struct NodeWgrad1Params {
  long long N;
  const __half* dz1;        // [N,128] fp16
  const float* x;
  const float* aggr;
  float* grad;
  int32_t* ready_in;        // or null: tile-level chaining with the node_bwd launch before this one (mlp_common.cuh):
  int32_t* sched;           // ready flags set by node_bwd, {next tile, CTAs done}
  long long* dbg;           // developer trace (MGN_CTA_TRACE builds)
  int dbg_cap;
};

__global__ void __launch_bounds__(NT, 1) node_wgrad1_kernel(const NodeWgrad1Params p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  check_smem_alignment(smem);
  Tail2* tail = reinterpret_cast<Tail2*>(smem + 6 * SLOT);
  uint64_t* bar_set = &tail->bar[0];    // [2]: MMAs reading operand set b have completed
  uint32_t& tmem_slot = tail->tmem_slot;
  const Map m = make_map();
  if (m.tid == 0) {
    mbar_init(&bar_set[0], 1);
    mbar_init(&bar_set[1], 1);
    mbar_init_fence();
  }
  if (m.warp == 0) tmem_alloc<256>(&tmem_slot);
  // x and aggr (forward pass) of the first tile into L2 while the previous kernel drains
  if (m.tid == 64) {
    prefetch_tile_rows_l2(p.x, (long long)blockIdx.x * TILE_ROWS, p.N, 512);
    prefetch_tile_rows_l2(p.aggr, (long long)blockIdx.x * TILE_ROWS, p.N, 512);
  }
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const bool chained = p.ready_in != nullptr;
  if (!chained) pdl_wait();     // the previous kernel in the stream is complete: global memory may be touched
  pdl_launch();                 // (common.cuh) the next kernel may start its own prologue under this one
  const uint32_t qoff = m.lane_off + m.c0;
  const uint32_t tm = tmem_slot;
  const long long num_tiles = (p.N + TILE_ROWS - 1) / TILE_ROWS;
  const uint64_t pol_keep = policy_evict_last();
  int* s_tile = reinterpret_cast<int*>(tail->red);
  uint32_t it = 0;
  cta_trace(p.dbg, p.dbg_cap, 1, 0);
  for (long long tile = blockIdx.x;; tile += gridDim.x, ++it) {
    if (it == 1) cta_trace(p.dbg, p.dbg_cap, 1, 1);
    if (chained) {      // next tile in index order = the order node_bwd finishes them in; wait for its rows
      if (m.tid == 0) {
        const int t = atomicAdd(p.sched, 1);
        if (t < num_tiles) tile_acquire(p.ready_in, t);
        *s_tile = t;
      }
      __syncthreads();
      tile = *s_tile;   // (thread 0 writes it again only after the barrier before the products below)
    }
    if (tile >= num_tiles) break;
    const long long row0 = tile * TILE_ROWS;
    const uint32_t b = it & 1;
    uint8_t* sG = smem + b * 3 * SLOT;          // 32 KB
    uint8_t* sU = sG + SLOT;                    // 128 x 256 fp16 = 64 KB
    auto rf = [&](int i) -> long long { const long long g = row0 + m.warp * 8 + i; return g < p.N ? g : -1; };
    if (it >= 2) mbar_wait(&bar_set[b], ((it >> 1) - 1) & 1);
    if (chained) load8_f16_to_tile_cg(sG, p.dz1, m, rf);
    else load8_f16_to_tile(sG, p.dz1, m, rf);
    load8_f32_to_tile(sU, 128, 0, p.x, m, rf, pol_keep);
    load8_f32_to_tile(sU, 128, 128, p.aggr, m, rf, pol_keep);
    if (m.tid == 32 && tile + gridDim.x < num_tiles) {      // the (likely) next tile's inputs into L2, a tile time ahead
      const long long n0 = (tile + gridDim.x) * TILE_ROWS;
      if (!chained) prefetch_tile_rows_l2(p.dz1, n0, p.N, 256);
      prefetch_tile_rows_l2(p.x, n0, p.N, 512);
      prefetch_tile_rows_l2(p.aggr, n0, p.N, 512);
    }
    fence_async_smem();
    fence_before_sync();
    __syncthreads();
    if (m.tid == 0) {
      fence_after_sync();
      const uint32_t uG = smem_u32(sG), uU = smem_u32(sU);
      const uint32_t idesc = make_idesc(128, 128, 1, 1);
      for (uint32_t ks = 0; ks < 8; ++ks)
        mma_f16(tm, desc_mnmajor(uG, 128, ks, 0), desc_mnmajor(uU, 128, ks, 0), idesc, (ks > 0 || it > 0) ? 1u : 0u);
      for (uint32_t ks = 0; ks < 8; ++ks)
        mma_f16(tm + 128, desc_mnmajor(uG, 128, ks, 0), desc_mnmajor(uU, 128, ks, 128), idesc,
                (ks > 0 || it > 0) ? 1u : 0u);
      mma_commit(&bar_set[b]);
    }
  }
  cta_trace(p.dbg, p.dbg_cap, 1, 2);
  if (chained) {
    if (m.tid == 0) sched_done(p.sched);
    pdl_wait();       // node_bwd is complete (its gradient flush included): what this launch's successors rely on
  }
  cta_trace(p.dbg, p.dbg_cap, 1, 3);
  // wait for the last commit of each set (commits complete in order; waiting both is simplest)
  if (it >= 1) mbar_wait(&bar_set[(it - 1) & 1], ((it - 1) >> 1) & 1);
  if (it >= 2) mbar_wait(&bar_set[it & 1], ((it - 2) >> 1) & 1);
  fence_after_sync();
  __syncthreads();          // every thread is past its last operand store: the operand buffers are dead
  if (it > 0) {             // (a chained CTA that arrived after the last tile was taken has nothing to add)
    flush_acc_bulk(tm + qoff, p.grad + GB_N_W1, 256, smem, m, policy_evict_last());
    flush_acc_bulk(tm + 128 + qoff, p.grad + GB_N_W1 + 128, 256, smem, m, policy_evict_last());
  }
  fence_before_sync();
  __syncthreads();
  if (m.warp == 0) tmem_dealloc<256>(tmem_slot);
}

// ===================================================================================== edge layer-1 wgrad
// dWe [128,128] += dz1^T e16 over all edge tiles, from the two fp16 tile-image streams (edge_bwd's dz1 images and the
// forward's images of e): both are already in the shared-memory operand layout, so a tile is two 32 KB bulk copies and
// eight MMAs.  Warp-specialised: thread 0 feeds a 3-stage ring, thread 32 issues the products (their completion
// releases the stage), everybody flushes the TMEM-resident accumulator at the end.
struct EdgeWgrad1Params {
  long long E;
  const __half* dz16;
  const __half* e16;
  float* grad;
};

__global__ void __launch_bounds__(NT, 1) edge_wgrad1_kernel(const EdgeWgrad1Params p) {
  constexpr int STAGES = 3;
  extern __shared__ __align__(1024) uint8_t smem[];
  check_smem_alignment(smem);
  Tail2* tail = reinterpret_cast<Tail2*>(smem + STAGES * 2 * SLOT);
  uint64_t* full = &tail->bar[0];      // [3] both tiles of the stage have landed
  uint64_t* empty = &tail->bar[3];     // [3] the products reading the stage have completed
  uint64_t& bar_done = tail->bar[6];
  uint32_t& tmem_slot = tail->tmem_slot;
  const Map m = make_map();
  if (m.tid == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    mbar_init(&bar_done, 1);
    mbar_init_fence();
  }
  if (m.warp == 0) tmem_alloc<128>(&tmem_slot);
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  pdl_wait();        // the previous kernel in the stream is complete: global memory may be touched from here on
  pdl_launch();      // (common.cuh) the next kernel may start its own prologue under this one
  const uint32_t tm = tmem_slot;
  const long long num_tiles = (p.E + TILE_ROWS - 1) / TILE_ROWS;
  const uint32_t my_tiles = blockIdx.x < num_tiles ? (uint32_t)((num_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x) : 0;
  if (m.tid == 0) {                       // producer
    for (uint32_t i = 0; i < my_tiles; ++i) {
      const uint32_t s = i % STAGES, n = i / STAGES;
      if (n > 0) mbar_wait(&empty[s], (n - 1) & 1);
      const size_t off = (size_t)(blockIdx.x + (long long)i * gridDim.x) * SLOT;
      mbar_arrive_expect_tx(&full[s], 2 * SLOT);
      bulk_g2s(smem + (2 * s) * SLOT, reinterpret_cast<const uint8_t*>(p.dz16) + off, SLOT, &full[s]);
      bulk_g2s(smem + (2 * s + 1) * SLOT, reinterpret_cast<const uint8_t*>(p.e16) + off, SLOT, &full[s]);
    }
  } else if (m.tid == 32) {               // MMA issuer
    for (uint32_t i = 0; i < my_tiles; ++i) {
      const uint32_t s = i % STAGES, n = i / STAGES;
      mbar_wait(&full[s], n & 1);
      fence_after_sync();
      umma_tile(tm, smem_u32(smem + (2 * s) * SLOT), 128, 1, smem_u32(smem + (2 * s + 1) * SLOT), 128, 1, 0, 128, 8,
                i > 0);
      mma_commit(&empty[s]);
    }
    mma_commit(&bar_done);
  }
  __syncwarp();
  if (my_tiles > 0) {
    mbar_wait(&bar_done, 0);
    fence_after_sync();
    flush_acc_bulk(tm + m.lane_off + m.c0, p.grad + GB_E_W1, 384, smem, m, policy_evict_last());
  }
  fence_before_sync();
  __syncthreads();
  if (m.warp == 0) tmem_dealloc<128>(tmem_slot);
}

// ===================================================================================== C ABI
static constexpr int EDGE_BWD_SMEM = 7 * SLOT + TAIL_BYTES;
static constexpr int NODE_BWD_SMEM = 6 * SLOT + TAIL_BYTES;
static constexpr int P_BWD_SMEM = 5 * SLOT + TAIL_BYTES;
static constexpr int NODE_WGRAD1_SMEM = 6 * SLOT + TAIL_BYTES;
static constexpr int EDGE_WGRAD1_SMEM = 6 * SLOT + TAIL_BYTES;

extern "C" int64_t mgn_grad_block_floats(void) { return GB_FLOATS; }

// Backward of the edge update + aggregation (autograd of layers.py:687-698, 703-719).
// dz16_out receives dz1 (the gradient at the first layer's pre-activation) as fp16 tile images for mgn_p_bwd.
extern "C" int mgn_edge_bwd(int64_t E, int64_t N, const float* e_in, const void* e16, const int32_t* rowmap,
                            const float* g_eout, const float* g_aggr, const void* P, const int32_t* senders,
                            const int32_t* receivers, float* g_e, void* dz16_out, const void* packed,
                            const float* b2, const float* b3, const float* b4, const float* scale,
                            float* grad_block, int norm_dim, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  MGN_CHECK_ARG(E >= 0 && N > 0, "edge_bwd: bad sizes");
  MGN_CHECK_ARG(norm_dim > 0 && norm_dim <= H, "edge_bwd: norm_dim must be in 1..128");
  if (E == 0) return MGN_OK;
  MGN_CHECK_ARG((e_in || e16) && g_aggr && P && senders && receivers && g_e && dz16_out && packed && b2 && b3 && b4 &&
                    scale && grad_block, "edge_bwd: null pointer");
  EdgeBwdParams p{};
  p.E = E; p.e_in = e_in; p.e16 = (const __half*)e16; p.rowmap = rowmap; p.g_eout = g_eout; p.g_aggr = g_aggr; p.P = (const __half*)P;
  p.senders = senders; p.receivers = receivers; p.g_e = g_e; p.dz16_out = (__half*)dz16_out;
  p.W = reinterpret_cast<const __half*>(reinterpret_cast<const uint8_t*>(packed) + PK_EDGE);
  p.b2 = b2; p.b3 = b3; p.b4 = b4; p.scale = scale; p.grad = grad_block;
  p.norm_d = (float)norm_dim; p.inv_sqrt_d = 1.0f / sqrtf((float)norm_dim);
  p.dbg = g_mgn_timeline_host; p.dbg_cap = g_mgn_timeline_cap_host;
  static const int dbg_skip = getenv("MGN_DEBUG_SKIP") ? atoi(getenv("MGN_DEBUG_SKIP")) : 0;
  p.dbg_skip = dbg_skip;
  static bool attr = false;
  if (!attr) {
    MGN_CUDA(cudaFuncSetAttribute(edge_bwd_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, EDGE_BWD_SMEM));
    MGN_CUDA(cudaFuncSetAttribute(edge_bwd_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, EDGE_BWD_SMEM));
    attr = true;
  }
  const int grid = (int)std::min<int64_t>(mgn_num_sms(), ceil_div64(E, TILE_ROWS));
  static const bool no_fast = getenv("MGN_EDGE_BWD_GENERAL") != nullptr;      // developer switch: always the general path
  if (!no_fast && rowmap == nullptr && g_eout != nullptr && e16 != nullptr)
    MGN_CUDA(mgn_launch(edge_bwd_kernel<true>, grid, NT, EDGE_BWD_SMEM, stream, p));
  else
    MGN_CUDA(mgn_launch(edge_bwd_kernel<false>, grid, NT, EDGE_BWD_SMEM, stream, p));
  MGN_LAUNCH_CHECK();
  return MGN_OK;
}

// mgn_edge_bwd for the processor loop when mgn_edge_fwd_hl saved the edge MLP's hidden activations: starts from the
// images h1, h2, h3 (no recompute of layers 1-3, no P, no e image).  g_e [E,128] holds the upstream edge gradient on
// entry (zeros if there is none) and g_e + dz1 We on exit (in place; edges in CSR order).  dz16_out as mgn_edge_bwd.
// Weight gradients of layers 2-4, their biases and the scale accumulate into grad_block; layer 1 is mgn_edge_wgrad1 +
// mgn_p_bwd as before.
extern "C" int mgn_edge_bwd_saved(int64_t E, int64_t N, const void* h1, const void* h2, const void* h3,
                                  const float* g_aggr, const int32_t* receivers, float* g_e, void* dz16_out,
                                  const void* packed, const float* b4, const float* scale, float* grad_block,
                                  int norm_dim, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  MGN_CHECK_ARG(E >= 0 && N > 0, "edge_bwd_saved: bad sizes");
  MGN_CHECK_ARG(norm_dim > 0 && norm_dim <= H, "edge_bwd_saved: norm_dim must be in 1..128");
  if (E == 0) return MGN_OK;
  MGN_CHECK_ARG(h1 && h2 && h3 && g_aggr && receivers && g_e && dz16_out && packed && b4 && scale && grad_block,
                "edge_bwd_saved: null pointer");
  EdgeBwdSavedParams p{};
  p.E = E; p.h1 = (const __half*)h1; p.h2 = (const __half*)h2; p.h3 = (const __half*)h3; p.g_aggr = g_aggr;
  p.receivers = receivers; p.g_e = g_e; p.dz16_out = (__half*)dz16_out;
  p.W = reinterpret_cast<const __half*>(reinterpret_cast<const uint8_t*>(packed) + PK_EDGE);
  p.b4 = b4; p.scale = scale; p.grad = grad_block;
  p.norm_d = (float)norm_dim; p.inv_sqrt_d = 1.0f / sqrtf((float)norm_dim);
  p.dbg = g_mgn_timeline_host; p.dbg_cap = g_mgn_timeline_cap_host;
  static bool attr = false;
  if (!attr) {
    MGN_CUDA(cudaFuncSetAttribute(edge_bwd_saved_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, EDGE_BWD_SMEM));
    attr = true;
  }
  const int grid = (int)std::min<int64_t>(mgn_num_sms(), ceil_div64(E, TILE_ROWS));
  MGN_CUDA(mgn_launch(edge_bwd_saved_kernel, grid, NT, EDGE_BWD_SMEM, stream, p));
  MGN_LAUNCH_CHECK();
  return MGN_OK;
}

// Backward of the node update (autograd of layers.py:733-746, 699) from the fp16 hidden saves.
// flags bit 0: the block has no RMSNorm (see mgn_node_fwd2)
extern "C" int mgn_node_bwd3(int64_t N, const float* g_xout, const void* h1, const void* h2, const void* h3,
                             float* g_x, float* g_aggr, void* dz1, const void* packed, const float* b4,
                             const float* scale, float* grad_block, int flags, int norm_dim, int32_t* ready_in,
                             int32_t* ready_out, void* stream_);
extern "C" int mgn_node_bwd2(int64_t N, const float* g_xout, const void* h1, const void* h2, const void* h3,
                             float* g_x, float* g_aggr, void* dz1, const void* packed, const float* b4,
                             const float* scale, float* grad_block, int flags, int norm_dim, void* stream_) {
  return mgn_node_bwd3(N, g_xout, h1, h2, h3, g_x, g_aggr, dz1, packed, b4, scale, grad_block, flags, norm_dim, nullptr,
                       nullptr, stream_);
}
extern "C" int mgn_node_bwd3(int64_t N, const float* g_xout, const void* h1, const void* h2, const void* h3,
                             float* g_x, float* g_aggr, void* dz1, const void* packed, const float* b4,
                             const float* scale, float* grad_block, int flags, int norm_dim, int32_t* ready_in,
                             int32_t* ready_out, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  MGN_CHECK_ARG(N > 0 && g_xout && h1 && h2 && h3 && g_x && g_aggr && dz1 && packed && b4 && scale && grad_block,
                "node_bwd: bad args");
  NodeBwdParams p{};
  p.N = N; p.g_xout = g_xout; p.h1 = (const __half*)h1; p.h2 = (const __half*)h2; p.h3 = (const __half*)h3;
  p.g_x = g_x; p.g_aggr = g_aggr; p.dz1 = (__half*)dz1;
  p.W = reinterpret_cast<const __half*>(reinterpret_cast<const uint8_t*>(packed) + PK_NODE);
  p.b4 = b4; p.scale = scale; p.grad = grad_block; p.no_norm = flags & 1;
  p.dbg = g_mgn_timeline_host; p.dbg_cap = g_mgn_timeline_cap_host;
  p.ready_out = ready_out;
  p.ready_in = ready_in;
  MGN_CHECK_ARG(norm_dim > 0 && norm_dim <= H, "node_bwd: norm_dim must be in 1..128");
  p.norm_d = (float)norm_dim; p.inv_sqrt_d = 1.0f / sqrtf((float)norm_dim);
  static bool attr = false;
  if (!attr) {
    MGN_CUDA(cudaFuncSetAttribute(node_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, NODE_BWD_SMEM));
    attr = true;
  }
  const int grid = (int)std::min<int64_t>(mgn_num_sms(), ceil_div64(N, TILE_ROWS));
  MGN_CUDA(mgn_launch(node_bwd_kernel, grid, NT, NODE_BWD_SMEM, stream, p));
  MGN_LAUNCH_CHECK();
  return MGN_OK;
}

extern "C" int mgn_node_bwd(int64_t N, const float* g_xout, const void* h1, const void* h2, const void* h3,
                            float* g_x, float* g_aggr, void* dz1, const void* packed, const float* b4,
                            const float* scale, float* grad_block, void* stream_) {
  return mgn_node_bwd2(N, g_xout, h1, h2, h3, g_x, g_aggr, dz1, packed, b4, scale, grad_block, 0, H, stream_);
}

// Backward of the gathers x[col], x[row]: dP [N,256] fp16 = [sum of dz1 over received edges | over sent edges] from
// edge_bwd's dz1 tile images, through the receiver CSR and the sender-sorted lists.  Fixed order, no atomics.
extern "C" int mgn_dp_gather(int64_t N, const void* dz16, const int32_t* rowptr, const int32_t* rowptr_s,
                             const int32_t* out_pos, void* dP, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  MGN_CHECK_ARG(N > 0 && dz16 && rowptr && rowptr_s && out_pos && dP, "dp_gather: bad args");
  DpGatherParams p{N, (const __half*)dz16, rowptr, rowptr_s, out_pos, (__half*)dP};
  const int64_t blocks = ceil_div64(2 * N * 16, 256);
  const int grid = (int)std::min<int64_t>(blocks, (int64_t)mgn_num_sms() * 8);
  MGN_CUDA(mgn_launch(dp_gather_kernel, grid, 256, 0, stream, p));
  MGN_LAUNCH_CHECK();
  return MGN_OK;
}

// Backward of the per-node partial products (mgn_p_fwd): g_x += dP [W_i ; W_j], dW_i, dW_j, db1.
extern "C" int mgn_p_bwd_c(int64_t N, const void* dP, const float* x, float* g_x, const void* packed,
                           float* grad_block, int32_t* ready_out, void* stream_);
extern "C" int mgn_p_bwd(int64_t N, const void* dP, const float* x, float* g_x, const void* packed,
                         float* grad_block, void* stream_) {
  return mgn_p_bwd_c(N, dP, x, g_x, packed, grad_block, nullptr, stream_);
}
extern "C" int mgn_p_bwd_c(int64_t N, const void* dP, const float* x, float* g_x, const void* packed,
                           float* grad_block, int32_t* ready_out, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  MGN_CHECK_ARG(N > 0 && dP && x && g_x && packed && grad_block, "p_bwd: bad args");
  PBwdParams p{N, (const __half*)dP, x, g_x,
               reinterpret_cast<const __half*>(reinterpret_cast<const uint8_t*>(packed) + PK_WCAT), grad_block, ready_out};
  static bool attr = false;
  if (!attr) {
    MGN_CUDA(cudaFuncSetAttribute(p_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, P_BWD_SMEM));
    attr = true;
  }
  const int grid = (int)std::min<int64_t>(mgn_num_sms(), ceil_div64(N, TILE_ROWS));
  MGN_CUDA(mgn_launch(p_bwd_kernel, grid, NT, P_BWD_SMEM, stream, p));
  MGN_LAUNCH_CHECK();
  return MGN_OK;
}

// Weight gradient of node_block layer 1: dW1' += dz1'^T [x | aggr].
extern "C" int mgn_node_wgrad1c(int64_t N, const void* dz1, const float* x, const float* aggr, float* grad_block,
                                int32_t* ready_in, int32_t* sched, void* stream_);
extern "C" int mgn_node_wgrad1(int64_t N, const void* dz1, const float* x, const float* aggr, float* grad_block,
                               void* stream_) {
  return mgn_node_wgrad1c(N, dz1, x, aggr, grad_block, nullptr, nullptr, stream_);
}
extern "C" int mgn_node_wgrad1c(int64_t N, const void* dz1, const float* x, const float* aggr, float* grad_block,
                                int32_t* ready_in, int32_t* sched, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  MGN_CHECK_ARG(N > 0 && dz1 && x && aggr && grad_block, "node_wgrad1: bad args");
  MGN_CHECK_ARG((ready_in == nullptr) == (sched == nullptr), "node_wgrad1: ready_in and sched go together");
  NodeWgrad1Params p{N, (const __half*)dz1, x, aggr, grad_block, ready_in, sched, g_mgn_timeline_host,
                     g_mgn_timeline_cap_host};
  static bool attr = false;
  if (!attr) {
    MGN_CUDA(cudaFuncSetAttribute(node_wgrad1_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, NODE_WGRAD1_SMEM));
    attr = true;
  }
  const int grid = (int)std::min<int64_t>(mgn_num_sms(), ceil_div64(N, TILE_ROWS));
  MGN_CUDA(mgn_launch(node_wgrad1_kernel, grid, NT, NODE_WGRAD1_SMEM, stream, p));
  MGN_LAUNCH_CHECK();
  return MGN_OK;
}

// Weight gradient of edge_block layer 1, edge-latent part: dWe += dz1^T e from the two tile-image streams
// (edge_block.0.weight[:, 0:128]).
extern "C" int mgn_edge_wgrad1(int64_t E, const void* dz16, const void* e16, float* grad_block, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  MGN_CHECK_ARG(E >= 0, "edge_wgrad1: bad size");
  if (E == 0) return MGN_OK;
  MGN_CHECK_ARG(dz16 && e16 && grad_block, "edge_wgrad1: null pointer");
  EdgeWgrad1Params p{E, (const __half*)dz16, (const __half*)e16, grad_block};
  static bool attr = false;
  if (!attr) {
    MGN_CUDA(cudaFuncSetAttribute(edge_wgrad1_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, EDGE_WGRAD1_SMEM));
    attr = true;
  }
  const int grid = (int)std::min<int64_t>(mgn_num_sms(), ceil_div64(E, TILE_ROWS));
  MGN_CUDA(mgn_launch(edge_wgrad1_kernel, grid, NT, EDGE_WGRAD1_SMEM, stream, p));
  MGN_LAUNCH_CHECK();
  return MGN_OK;
}
