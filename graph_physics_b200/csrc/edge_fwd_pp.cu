// edge_fwd_pp.cu -- the edge forward kernel of steps 1..K-1 (operand tile images available), with TWO tiles in flight
// per CTA so that no thread ever waits for a tensor-core product.
//
// Same arithmetic as mlp_fwd_kernel<true> in mlp_fwd.cu (GraphNetBlock.forward, graphphysics/models/layers.py:687-698:
// edge_block on [e | x[col] | x[row]], RMSNorm, residual, sum at the receivers); what changes is the schedule.  A tile
// goes through six stages
//   O0 issue product 1 (e16 image x We)            O3 epilogue 3, issue product 4
//   O1 epilogue 1 (+ P_i + P_j, ReLU), issue 2     O4 bias + RMSNorm of product 4 (registers)
//   O2 epilogue 2, issue product 3                 O5 residual add, stores, image of the output rows, segmented sum
// and iteration k of the main loop runs  O3(k-1) O1(k) O4(k-1) O2(k) O5(k-1) O0(k+1): every product is issued one
// stage before the OTHER tile's epilogue starts and is waited for one stage after it ends, so the tensor pipe works
// under the epilogues and the 1-2 k cycles of issue-to-commit latency per product disappear from the critical path.
//
// Resources.  Tensor memory: two slots of [128 accumulator columns | 64 columns holding the fp16 hidden activations as
// the next product's A operand] (the activations never touch shared memory).  Shared memory: the four weight tiles
// (resident), ONE operand-image buffer (free again as soon as product 1 of a tile is done; the next image arrives
// three stages before its use), ONE P_j staging buffer, and a 32 KB fp32 staging buffer through which the output rows
// pass in two halves of 64 rows (row-owner registers -> coalesced rows).
#include <algorithm>

#include "mlp_v2.cuh"
#include "mlp_fwd_abi.cuh"

using namespace mgn;

struct EdgePPParams {
  long long E;
  const float* e_in;        // [E,128] fp32, CSR order: residual stream
  const __half* e16_in;     // tile images of e_in (written by the previous step)
  float* e_out;             // [E,128] fp32, CSR order
  __half* e16_next;         // tile images of e_out, or null
  const __half* P;          // [N,256] fp16
  const int32_t* senders;
  const int32_t* receivers;
  float* aggr;              // [N,128] fp32, zero on entry
  const __half* W;          // We, W2, W3, W4 tiles
  const float* b2;
  const float* b3;
  const float* b4;
  const float* scale;
  long long* dbg;
  int dbg_cap;
};

__global__ void __launch_bounds__(NT, 1) edge_fwd_pp_kernel(const EdgePPParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  check_smem_alignment(smem);
  uint8_t* sW = smem;                          // 4 x 32 KB
  uint8_t* sA = smem + 4 * TILE_BYTES;         // operand image
  uint8_t* sPj = smem + 5 * TILE_BYTES;        // staged P_j rows (fp16)
  uint8_t* stage = smem + 6 * TILE_BYTES;      // fp32 staging, 64 rows
  Tail2* tail = reinterpret_cast<Tail2*>(smem + 7 * TILE_BYTES);
  uint64_t& wbar = tail->bar[0];
  uint64_t* mbar = &tail->bar[1];              // [2]
  uint64_t& bar_in = tail->bar[3];
  uint32_t& tmem_slot = tail->tmem_slot;
  const Map m = make_map();

  if (m.tid == 0) {
    mbar_init(&wbar, 1);
    mbar_init(&mbar[0], 1);
    mbar_init(&mbar[1], 1);
    mbar_init(&bar_in, 1);
    mbar_init_fence();
  }
  if (m.warp == 0) tmem_alloc<512>(&tmem_slot);
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  pdl_wait();        // the previous kernel in the stream is complete: global memory may be touched from here on
  pdl_launch();      // (common.cuh) the next kernel may start its own prologue under this one
  const uint32_t tm = tmem_slot;
  const uint32_t sA_u32 = smem_u32(sA), sW_u32 = smem_u32(sW), sPj_u32 = smem_u32(sPj);
  if (m.tid == 0) {
    mbar_arrive_expect_tx(&wbar, 4 * TILE_BYTES);
    for (uint32_t o = 0; o < 4 * TILE_BYTES; o += TILE_BYTES)
      bulk_g2s(sW + o, reinterpret_cast<const uint8_t*>(p.W) + o, TILE_BYTES, &wbar);
  }
  // b2, b3, b4, scale live in shared memory (the tail's 2 KB scratch): with the products already complete when an
  // epilogue starts, a global load of its bias would be the longest wait of the stage.  The row-sum scratch of the
  // RMSNorm moves into the staging buffer, which is idle outside O5.
  float* svec = tail->red;                     // [4][128]
  float* red = reinterpret_cast<float*>(stage);
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
  const uint64_t pol_stream = policy_evict_first();
  const uint64_t pol_keep = policy_evict_last();
  const long long num_tiles = (p.E + TILE_ROWS - 1) / TILE_ROWS;
  const long long first = blockIdx.x, stride = gridDim.x;
  const int K = first < num_tiles ? (int)((num_tiles - first + stride - 1) / stride) : 0;
  auto tile_of = [&](int k) -> long long { return first + (long long)k * stride; };
  // coalesced mapping: lane j < 8 of warp w stands for tile row 8w + j (as in mlp_fwd.cu).  The output passes through
  // the staging buffer in two halves: rows 8w..8w+3 of every warp, then rows 8w+4..8w+7, so that a warp still sums its
  // eight consecutive rows into one partial per receiver (same aggregation order as the one-tile kernel)
  auto crow = [&](int j) -> int { return 8 * m.warp + j; };

  struct Idx { int rcv_o, dst_c, rcv_c, snd_c; };
  auto load_idx = [&](int k) -> Idx {
    Idx ix{0, -1, -1, -1};
    if (k >= K) return ix;
    const long long r0 = tile_of(k) * TILE_ROWS;
    if (r0 + m.row < p.E) ix.rcv_o = __ldg(p.receivers + r0 + m.row);
    const long long g = r0 + crow(m.lane & 7);
    if (g < p.E) {
      ix.dst_c = (int)(g - r0);          // row inside the tile; the global row is r0 + dst_c
      ix.rcv_c = __ldg(p.receivers + g);
      ix.snd_c = __ldg(p.senders + g);
    }
    return ix;
  };
  uint32_t img_n = 0;                          // thread 0: images waited for so far
  auto issue_img = [&](int k) {                // thread 0
    mbar_arrive_expect_tx(&bar_in, TILE_BYTES);
    bulk_g2s(sA, reinterpret_cast<const uint8_t*>(p.e16_in) + (size_t)tile_of(k) * TILE_BYTES, TILE_BYTES, &bar_in);
  };
  const int hw = m.lane >> 4, q16 = m.lane & 15;
  auto issue_pj = [&](const Idx& ix) {         // all threads: half a warp per 256-byte row
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int j = 2 * i + hw;
      const int rc = __shfl_sync(0xffffffffu, ix.rcv_c, j);
      const int sn = __shfl_sync(0xffffffffu, ix.snd_c, j);
      cp_async16(sPj_u32 + tile_off(128, crow(j), 8 * q16),
                 p.P + (long long)(rc >= 0 ? sn : 0) * 256 + 128 + 8 * q16, rc >= 0 ? 16u : 0u, pol_keep);
    }
  };
  uint32_t ph[2] = {0u, 0u};
  auto acc_of = [&](int s) -> uint32_t { return tm + 192u * (uint32_t)s; };
  auto aop_of = [&](int s) -> uint32_t { return tm + 192u * (uint32_t)s + 128u; };
  auto mma_wait = [&](int s) {
    mbar_wait(&mbar[s], ph[s]);
    ph[s] ^= 1u;
    fence_after_sync();
  };
  auto acts_ready = [&]() {
    tmem_st_wait();
    fence_before_sync();
    __syncthreads();
  };
  Timeline tl;
  tl.init(m.tid == 0 && blockIdx.x == 0, p.dbg, p.dbg_cap);

  // ------------------------------------------------------------------ stages
  auto O0 = [&](int k) {                       // product 1 of tile k
    if (m.tid == 0) {
      mbar_wait(&bar_in, img_n & 1u);
      ++img_n;
      fence_after_sync();
      umma_tile(acc_of(k & 1), sA_u32, 128, 0, sW_u32, 128, 0, 0, 128, 8, 0);
      mma_commit(&mbar[k & 1]);
    }
  };
  // P_i[receiver] chunk of this thread's row: requested one stage before O1 needs it (registers)
  uint4 pi[4];
  auto load_pi = [&](int k, const Idx& ix) {
    const bool ok = k < K && tile_of(k) * TILE_ROWS + m.row < p.E;
    const __half* Pi = p.P + (long long)(ok ? ix.rcv_o : 0) * 256 + m.c0;
#pragma unroll
    for (int j = 0; j < 4; ++j) pi[j] = ok ? ld_u4_hint(Pi + 8 * j, pol_keep) : make_uint4(0, 0, 0, 0);
  };
  auto O1 = [&](int k, const Idx& ix, const Idx& ixn) {
    const int s = k & 1;
    cp_async_wait_all();
    __syncthreads();                           // the staged P_j rows of tile k are visible
    mma_wait(s);
    if (m.tid == 0 && k + 1 < K) issue_img(k + 1);      // the image buffer is free: next tile's operand
    float v[32], b[32];
    tmem_ld32(acc_of(s) + m.lane_off + m.c0, v);
    load_row_chunk_f16(sPj, 128, m.row, m.c0, b);
    tmem_ld_wait();
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float a[8];
      unpack8(pi[j], a);
#pragma unroll
      for (int i = 0; i < 8; ++i) v[8 * j + i] = fmaxf(v[8 * j + i] + a[i] + b[8 * j + i], 0.f);
    }
    store_row_chunk_tmem(aop_of(s) + m.lane_off + (m.c0 >> 1), v, nullptr, 0, m.c0);
    acts_ready();                              // (also: every thread has read its P_j row)
    if (k + 1 < K) issue_pj(ixn);
    if (m.tid == 0) {
      fence_after_sync();
      umma_tile_ts(acc_of(s), aop_of(s), sW_u32 + TILE_BYTES, 128, 0, 0, 128, 8, 0);
      mma_commit(&mbar[s]);
    }
  };
  auto Omid = [&](int k, int layer) {           // layer 2 or 3: epilogue, then issue the next product
    const int s = k & 1;
    float4 bb[8];
    vec32(layer == 2 ? 0 : 1, bb);
    mma_wait(s);
    epi_bias_relu_tmem_q(acc_of(s) + m.lane_off + m.c0, aop_of(s) + m.lane_off + (m.c0 >> 1), bb, m, nullptr, 0);
    acts_ready();
    if (m.tid == 0) {
      fence_after_sync();
      umma_tile_ts(acc_of(s), aop_of(s), sW_u32 + (uint32_t)layer * TILE_BYTES, 128, 0, 0, 128, 8, 0);
      mma_commit(&mbar[s]);
    }
  };
  // O4 + O5 share registers (the normalised row chunk y and the residual rows), so they are one lambda with the other
  // tile's O2 passed in between as `mid`
  auto O45 = [&](int k, const Idx& ix, auto&& mid) {
    const int s = k & 1;
    const long long tile = tile_of(k);
    const long long r0 = tile * TILE_ROWS;
    {
      float v[32];
      float4 bb[8];
      vec32(2, bb);
      mma_wait(s);
      tmem_ld32(acc_of(s) + m.lane_off + m.c0, v);
      tmem_ld_wait();
      float ss = 0.f;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        v[4 * j] += bb[j].x; v[4 * j + 1] += bb[j].y; v[4 * j + 2] += bb[j].z; v[4 * j + 3] += bb[j].w;
        ss += v[4 * j] * v[4 * j] + v[4 * j + 1] * v[4 * j + 1] + v[4 * j + 2] * v[4 * j + 2] + v[4 * j + 3] * v[4 * j + 3];
      }
      fence_before_sync();                       // the accumulator slot is read: a later product may overwrite it
      ss = row_sum4(ss, red, m);
      const float inv = 1.0f / (sqrtf(ss) * INV_SQRT_H + RMS_EPS);
      float4 sc[8];
      vec32(3, sc);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        v[4 * j + 0] = sc[j].x * (v[4 * j + 0] * inv);
        v[4 * j + 1] = sc[j].y * (v[4 * j + 1] * inv);
        v[4 * j + 2] = sc[j].z * (v[4 * j + 2] * inv);
        v[4 * j + 3] = sc[j].w * (v[4 * j + 3] * inv);
      }
      // the normalised row chunk waits in the (dead) accumulator columns while the other tile's O2 runs: 32 registers
      // less across that stage
      tmem_st32(acc_of(s) + m.lane_off + m.c0, v);
      tmem_st_wait();
    }
    // residual rows (L2-prefetched two stages ago), coalesced mapping: the first half's four rows are requested before
    // the other tile's O2, the second half's after it (registers)
    float4 old[8];
    int dst[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) dst[i] = __shfl_sync(0xffffffffu, ix.dst_c, i);
#pragma unroll
    for (int i = 0; i < 4; ++i)
      old[i] = dst[i] >= 0 ? ld_f4_hint(p.e_in + (r0 + dst[i]) * H + 4 * m.lane, pol_stream)
                           : make_float4(0.f, 0.f, 0.f, 0.f);
    mid();
    // ---- O5: two halves of 64 rows through the fp32 staging buffer
#pragma unroll
    for (int i = 4; i < 8; ++i)
      old[i] = dst[i] >= 0 ? ld_f4_hint(p.e_in + (r0 + dst[i]) * H + 4 * m.lane, pol_stream)
                           : make_float4(0.f, 0.f, 0.f, 0.f);
    float v[32];
    tmem_ld32(acc_of(s) + m.lane_off + m.c0, v);
    tmem_ld_wait();
    fence_before_sync();                         // (the barriers below order this read before the slot's next product)
    int cur = -1;
    float4 sum = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      if (((m.row >> 2) & 1) == h) {
        const uint32_t srow = (uint32_t)((m.row >> 3) << 2) | (uint32_t)(m.row & 3);
#pragma unroll
        for (int j = 0; j < 8; ++j)
          *reinterpret_cast<float4*>(stage + stage_off(srow, m.c0 + 4 * j)) =
              make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
      }
      __syncthreads();
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int j = 4 * h + i;
        const int trow = 8 * m.warp + j;
        const float4 y = *reinterpret_cast<const float4*>(stage + stage_off(4 * m.warp + i, 4 * m.lane));
        const int rc = __shfl_sync(0xffffffffu, ix.rcv_c, j);
        uint8_t* img = p.e16_next ? reinterpret_cast<uint8_t*>(p.e16_next) + (size_t)tile * TILE_BYTES +
                                        tile_off(128, trow, 4 * m.lane)
                                  : nullptr;
        if (dst[j] < 0) {
          if (img) *reinterpret_cast<uint2*>(img) = make_uint2(0u, 0u);     // rows past the end: zeros in the image
          continue;
        }
        const float4 o = make_float4(old[j].x + y.x, old[j].y + y.y, old[j].z + y.z, old[j].w + y.w);
        st_f4_hint(p.e_out + (r0 + trow) * H + 4 * m.lane, o, pol_stream);
        if (img) {
          uint2 pk;
          pk.x = pack_h2(o.x, o.y);
          pk.y = pack_h2(o.z, o.w);
          *reinterpret_cast<uint2*>(img) = pk;
        }
        if (rc != cur) {
          if (cur >= 0) red_add_f4_hint(p.aggr + (long long)cur * H + 4 * m.lane, sum, pol_keep);
          cur = rc;
          sum = y;
        } else {
          sum.x += y.x; sum.y += y.y; sum.z += y.z; sum.w += y.w;
        }
      }
      __syncthreads();                           // staging buffer free
    }
    if (cur >= 0) red_add_f4_hint(p.aggr + (long long)cur * H + 4 * m.lane, sum, pol_keep);
  };
  auto prefetch_next = [&](int k, const Idx& ixn) {      // L2 prefetch of tile k's streamed / gathered rows
    if (k >= K) return;
    const long long g = tile_of(k) * TILE_ROWS + m.row;
    if (g < p.E) {
      prefetch_l2(reinterpret_cast<const char*>(p.e_in + g * H) + 128 * m.cq);
      if (m.cq < 2) prefetch_l2(reinterpret_cast<const char*>(p.P + (long long)ixn.rcv_o * 256) + 128 * m.cq);
    }
  };

  // ------------------------------------------------------------------ pipeline
  Idx ix_prev{0, -1, -1, -1};
  Idx ix_cur = load_idx(0);
  Idx ix_nxt = load_idx(1);
  if (K > 0) {
    if (m.tid == 0) {
      issue_img(0);
      mbar_wait(&wbar, 0);
    }
    issue_pj(ix_cur);
    load_pi(0, ix_cur);
    O0(0);
  }
  for (int k = 0; k <= K && K > 0; ++k) {
    tl.stamp();
    const bool has_prev = k >= 1, has_cur = k < K;
    if (has_prev) Omid(k - 1, 3);
    tl.stamp();
    if (has_cur) O1(k, ix_cur, ix_nxt);
    tl.stamp();
    if (has_prev) {
      O45(k - 1, ix_prev, [&]() {
        tl.stamp();
        if (has_cur) {
          Omid(k, 2);
          prefetch_next(k + 1, ix_nxt);
        }
        tl.stamp();
      });
    } else if (has_cur) {
      tl.stamp();
      Omid(k, 2);
      prefetch_next(k + 1, ix_nxt);
      tl.stamp();
    }
    tl.stamp();
    if (k + 1 < K) {
      load_pi(k + 1, ix_nxt);
      O0(k + 1);
    }
    ix_prev = ix_cur;
    ix_cur = ix_nxt;
    ix_nxt = load_idx(k + 2);
  }
  if (K == 0 && m.tid == 0) mbar_wait(&wbar, 0);    // never exit with the weight copy in flight

  fence_before_sync();
  __syncthreads();
  if (m.warp == 0) tmem_dealloc<512>(tmem_slot);
}

static constexpr int EDGE_PP_SMEM = 7 * 32768 + TAIL_BYTES;

// Steps 1..K-1 of the processor loop: e_in / e_out in CSR order, operand tile images of e_in available.
// Same results as mgn_edge_fwd(..., rowmap = NULL, e16_in = images) up to the order of the segmented-sum partials.
extern "C" int mgn_edge_fwd_pp(int64_t E, int64_t N, const float* e_in, const void* e16_in, float* e_out,
                               void* e16_next, const void* P, const int32_t* senders, const int32_t* receivers,
                               float* aggr, const void* packed, const float* b2, const float* b3, const float* b4,
                               const float* scale, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  MGN_CHECK_ARG(E >= 0 && N > 0 && aggr, "edge_fwd_pp: bad sizes");
  MGN_CUDA(cudaMemsetAsync(aggr, 0, sizeof(float) * N * H, stream));
  if (E == 0) return MGN_OK;
  MGN_CHECK_ARG(e_in && e16_in && e_out && P && senders && receivers && packed && b2 && b3 && b4 && scale,
                "edge_fwd_pp: null pointer");
  MGN_CHECK_ARG(e_in != e_out, "edge_fwd_pp: in-place not supported");
  EdgePPParams p{};
  p.E = E; p.e_in = e_in; p.e16_in = (const __half*)e16_in; p.e_out = e_out; p.e16_next = (__half*)e16_next;
  p.P = (const __half*)P; p.senders = senders; p.receivers = receivers; p.aggr = aggr;
  p.W = reinterpret_cast<const __half*>(reinterpret_cast<const uint8_t*>(packed) + PK_EDGE);
  p.b2 = b2; p.b3 = b3; p.b4 = b4; p.scale = scale;
  p.dbg = g_mgn_timeline_host; p.dbg_cap = g_mgn_timeline_cap_host;
  static bool attr = false;
  if (!attr) {
    MGN_CUDA(cudaFuncSetAttribute(edge_fwd_pp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, EDGE_PP_SMEM));
    attr = true;
  }
  const int64_t tiles = ceil_div64(E, TILE_ROWS);
  const int grid = (int)std::min<int64_t>(mgn_num_sms(), tiles);
  MGN_CUDA(mgn_launch(edge_fwd_pp_kernel, grid, NT, EDGE_PP_SMEM, stream, p));
  MGN_LAUNCH_CHECK();
  return MGN_OK;
}
