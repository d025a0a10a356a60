// mlp_fwd.cu -- fused forward kernels of one MeshGraphNet message-passing step.
//
// Replaces, per GraphNetBlock.forward (graphphysics/models/layers.py:667-701):
//   edge kernel: x[col]/x[row] gathers (689-690), cat (717), edge_block = 4 x Linear +
//                ReLU + RMSNorm (718, build_mlp 101-113), scatter-sum at edge_index[1]
//                (694-696) and the edge residual (698);
//   node kernel: cat[x, aggr] (744), node_block (745) and the node residual (699).
//
// Edge layer 1 uses the split W1 = [W_e | W_i | W_j]: the per-node products
// P = x W_i^T + b1 | x W_j^T are computed once per node (p kernel below) and gathered
// as fp16 rows in the layer-1 epilogue, so the per-edge tensor work is e W_e^T only.
// All products run on tcgen05 (fp16 operands, fp32 accumulation in TMEM); residual
// streams, RMSNorm and the segmented sum are fp32.
#define MGN_FILE_ID 1
#include <algorithm>

#include "mlp_v2.cuh"
#include "mlp_fwd_abi.cuh"

using namespace mgn;

struct FwdParams {
  long long M;                // rows: E (edge) or N (node)
  const float* in0;           // edge: e [E,128]; node: x [N,128]
  const float* in1;           // node: aggr [N,128]
  const int32_t* rowmap;      // edge: CSR position -> caller row of e (null = identity)
  int out_csr;                // edge: 1 = rowmap applies to the input rows only, the output is written in CSR order
  float* out;                 // e_out / x_out, same row convention as in0
  const __half* P;            // edge: [N,256] fp16 (P_i + b1 | P_j)
  const int32_t* senders;     // edge, CSR order
  const int32_t* receivers;   // edge, CSR order
  float* aggr;                // edge: [N,128] fp32, zero on entry
  float* carry;               // edge: [ceil(E/8),128] fp32 (deterministic aggregation: edge_fwd_hl.cu
                              // aggr_fixup_kernel, slots of 8 rows = one warp's rows here)
  float inv_sqrt_d;           // RMSNorm: d^-1/2 of the TRUE hidden size (zero-padded narrower models)
  const __half* W;            // packed fp16 weight tiles
  const float* b1;            // node only
  const float* b2;
  const float* b3;
  const float* b4;
  const float* scale;
  __half* save_h1;            // optional fp16 saves of the hidden activations.  edge: [M,128] rows (row = CSR position);
  __half* save_h2;            // node: tile images (32 KB swizzled 128 x 128 tile per 128 rows, bulk-stored as the tile
  __half* save_h3;            // sits in shared memory -- exactly what node_bwd's operand slots want)
  __half* e16_out;            // edge, optional: fp16 tile images of in0 (32 KB per 128-row tile) for edge_bwd
  const __half* e16_in;       // edge, optional: tile images of in0 made by the previous step (then in0 is only read
                              // for the residual and the operand tile arrives by bulk copy, one tile ahead)
  __half* e16_next;           // edge, optional: tile images of the OUTPUT rows (the next step's e16_in)
  __half* e16lo_next;         // edge, optional: tile images of the fp16 remainder  out - fp16(out)  (edge_fwd_hl.cu reads
                              // the residual stream as image + remainder); `out` may then be null
  long long* dbg;             // developer timeline (null = off)
  int dbg_cap;
  int32_t* ready_out;         // node update, or null: ready_out[tile] = 1 once the tile's output rows are written (chaining)
  int no_norm;                // node: the block has no RMSNorm (the decoder, build_mlp(..., layer_norm=False)): out = in0 + z4
};

template <bool EDGE>
__global__ void __launch_bounds__(NT, 1) mlp_fwd_kernel(const FwdParams p) {
  constexpr uint32_t W_BYTES = EDGE ? 4 * TILE_BYTES : 5 * TILE_BYTES;
  constexpr uint32_t A_BYTES = EDGE ? TILE_BYTES : 2 * TILE_BYTES;
  constexpr uint32_t K0_STEPS = EDGE ? 8 : 16;
  constexpr uint32_t W1_BYTES = EDGE ? TILE_BYTES : 2 * TILE_BYTES;

  extern __shared__ __align__(1024) uint8_t smem[];
  check_smem_alignment(smem);
  uint8_t* sW = smem;
  uint8_t* sA = smem + W_BYTES;
  // fp32 staging of the block output: its own 64 KB for the edge kernel, the (dead) operand tile for the node kernel
  uint8_t* stage = EDGE ? smem + W_BYTES + A_BYTES : sA;
  Tail2* tail = reinterpret_cast<Tail2*>(smem + W_BYTES + A_BYTES + (EDGE ? 2 * TILE_BYTES : 0));
  uint64_t& wbar = tail->bar[0];
  uint64_t& mbar = tail->bar[1];
  uint64_t& bar_in = tail->bar[2];
  uint32_t& tmem_slot = tail->tmem_slot;
  const Map m = make_map();
  const bool img_in = EDGE && p.e16_in != nullptr;

  if (m.tid == 0) {
    mbar_init(&bar_in, 1);
    mbar_init(&wbar, 1);
    mbar_init(&mbar, 1);
    mbar_init_fence();
  }
  // edge kernel: the hidden activations never touch shared memory -- each epilogue writes its fp16 row chunk into 64
  // TMEM columns next to the accumulator and the next product reads its A operand from there
  constexpr uint32_t TMEM_COLS = EDGE ? 256 : 128;
  if (m.warp == 0) tmem_alloc<TMEM_COLS>(&tmem_slot);
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  pdl_wait();        // the previous kernel in the stream is complete: global memory may be touched from here on
  pdl_launch();      // (common.cuh) the next kernel may start its own prologue under this one
  const uint32_t tmem_acc = tmem_slot;
  const uint32_t taddr = tmem_acc + m.lane_off + m.c0;
  const uint32_t tmem_a = tmem_acc + 128;
  const uint32_t taddr_a = tmem_a + m.lane_off + (m.c0 >> 1);

  if (m.tid == 0) {
    mbar_arrive_expect_tx(&wbar, W_BYTES);
    for (uint32_t o = 0; o < W_BYTES; o += TILE_BYTES)
      bulk_g2s(sW + o, reinterpret_cast<const uint8_t*>(p.W) + o, TILE_BYTES, &wbar);
  }

  const long long num_tiles = (p.M + TILE_ROWS - 1) / TILE_ROWS;
  uint32_t phase = 0;
  const uint32_t sA_u32 = smem_u32(sA), sW_u32 = smem_u32(sW);
  bool w_ready = false;
  auto mma_wait = [&]() {
    mbar_wait(&mbar, phase);
    phase ^= 1;
    fence_after_sync();
  };
  auto tile_ready = [&]() {
    fence_async_smem();
    fence_before_sync();
    __syncthreads();
  };
  auto acts_ready = [&]() {       // the TMEM-resident activations of every warp are written
    tmem_st_wait();
    fence_before_sync();
    __syncthreads();
  };

  const uint64_t pol_stream = policy_evict_first();   // edge / node latent streams: touched once or twice
  const uint64_t pol_keep = policy_evict_last();      // per-node arrays gathered by every edge
  // Index registers run one tile ahead (the loads for tile i+1 are issued during tile i), so no
  // thread waits on an index before it can issue its tile's row loads.
  struct Idx { int rcv_o, snd_o, dst_c, rcv_c, snd_c, prev_c; };   // prev_c: receiver of the row before this warp's 8 rows
  auto load_idx = [&](long long tile_) -> Idx {
    Idx ix{0, 0, -1, -1, -1, -1};
    if (tile_ >= num_tiles) return ix;
    const long long r0 = tile_ * TILE_ROWS;
    const long long s0 = r0 + m.warp * 8;
    if (EDGE && s0 > 0 && s0 < p.M) ix.prev_c = __ldg(p.receivers + s0 - 1);
    if (EDGE && r0 + m.row < p.M) {
      ix.rcv_o = __ldg(p.receivers + r0 + m.row);
      ix.snd_o = __ldg(p.senders + r0 + m.row);
    }
    const long long g = r0 + m.warp * 8 + (m.lane & 7);
    if (g < p.M) {
      ix.dst_c = (EDGE && p.rowmap) ? __ldg(p.rowmap + g) : (int)g;
      if (EDGE) {
        ix.rcv_c = __ldg(p.receivers + g);
        ix.snd_c = __ldg(p.senders + g);
      }
    }
    return ix;
  };
  Idx nxt = load_idx(blockIdx.x);
  auto issue_image = [&](long long tile_) {      // thread 0: operand tile of `tile_` by bulk copy
    mbar_arrive_expect_tx(&bar_in, TILE_BYTES);
    bulk_g2s(sA, reinterpret_cast<const uint8_t*>(p.e16_in) + (size_t)tile_ * TILE_BYTES, TILE_BYTES, &bar_in);
  };
  if (img_in && m.tid == 0 && (long long)blockIdx.x < num_tiles) issue_image(blockIdx.x);
  uint32_t it = 0;
  Timeline tl;
  tl.init(m.tid == 0 && blockIdx.x == 0, p.dbg, p.dbg_cap);

  for (long long tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
    tl.stamp();   // 0 tile start
    const long long row0 = tile * TILE_ROWS;
    const long long grow = row0 + m.row;          // epilogue (row-owner) mapping
    const bool rvalid = grow < p.M;
    const Idx cur = nxt;
    const int rcv_o = cur.rcv_o, snd_o = cur.snd_o, dst_c = cur.dst_c, rcv_c = cur.rcv_c;
    const int cur_idx_prev = cur.prev_c;

    // P_j[sender] rows (a true gather): asynchronous copies into the idle staging area, half a warp per 256-byte
    // row; they land while the operand tile is loaded and the layer-1 product runs.  P_i[receiver] rows are read by
    // the epilogue threads directly (consecutive tile rows share their receiver).
    if (EDGE) {
      const int hw = m.lane >> 4, q = m.lane & 15;
      const uint32_t st_u32 = smem_u32(stage);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int rc = __shfl_sync(0xffffffffu, cur.rcv_c, 2 * i + hw);
        const int sn = __shfl_sync(0xffffffffu, cur.snd_c, 2 * i + hw);
        cp_async16(st_u32 + tile_off(128, m.warp * 8 + 2 * i + hw, 8 * q),
                   p.P + (long long)(rc >= 0 ? sn : 0) * 256 + 128 + 8 * q, rc >= 0 ? 16u : 0u, pol_keep);
      }
    }
    // ---- load phase: fp32 rows -> fp16 operand tile (skipped when the previous step left the tile image)
    if (!img_in) {
      auto rf = [&](int i) -> long long { return (long long)__shfl_sync(0xffffffffu, dst_c, i); };
      load8_f32_to_tile(sA, 128, 0, p.in0, m, rf, pol_stream);
      if (!EDGE) load8_f32_to_tile(sA, 128, 128, p.in1, m, rf, pol_stream);
    }
    nxt = load_idx(tile + gridDim.x);
    if (!img_in) tile_ready();
    if (!w_ready) {
      mbar_wait(&wbar, 0);
      w_ready = true;
    }
    tl.stamp();   // 1 operand tile ready

    // ---- layer 1
    if (m.tid == 0) {
      if (img_in) mbar_wait(&bar_in, it & 1);
      fence_after_sync();
      const bool image = EDGE && p.e16_out != nullptr && !img_in;
      if (image) {      // the fp16 operand tile as it sits in shared memory is what edge_bwd bulk-loads
        bulk_s2g(reinterpret_cast<uint8_t*>(p.e16_out) + (size_t)tile * TILE_BYTES, sA, TILE_BYTES, pol_stream);
        bulk_commit();
      }
      umma_tile(tmem_acc, sA_u32, 128, 0, sW_u32, 128, 0, 0, 128, K0_STEPS, 0);
      if (image) bulk_wait_read0();   // the epilogue overwrites sA: commit only after the copy has read it
      mma_commit(&mbar);
    }
    if (EDGE) {
      uint4 pi[4];
      if (rvalid) {
        const __half* Pi = p.P + (long long)rcv_o * 256 + m.c0;
#pragma unroll
        for (int j = 0; j < 4; ++j) pi[j] = ld_u4_hint(Pi + 8 * j, pol_keep);
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j) pi[j] = make_uint4(0, 0, 0, 0);
      }
      cp_async_wait_all();
      __syncthreads();      // the staged P_j rows are visible to every thread
      mma_wait();
      tl.stamp();   // 2 layer-1 product done
      float v[32], b[32];
      tmem_ld32(taddr, v);
      load_row_chunk_f16(stage, 128, m.row, m.c0, b);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float a[8];
        unpack8(pi[j], a);
#pragma unroll
        for (int i = 0; i < 8; ++i) v[8 * j + i] = fmaxf(v[8 * j + i] + a[i] + b[8 * j + i], 0.f);
      }
      store_row_chunk_tmem(taddr_a, v, rvalid ? p.save_h1 : nullptr, grow, m.c0);
    } else {
      float4 bb[8];
      load_vec32(p.b1, m.c0, bb);
      mma_wait();
      epi_bias_relu_q(taddr, bb, sA, m, nullptr, 0);
    }
    if (EDGE) {
      // sA is dead from here on: the next tile's operand image can come in now, three products ahead of its use
      if (img_in && m.tid == 0 && tile + gridDim.x < num_tiles) issue_image(tile + gridDim.x);
      acts_ready();
    } else {
      tile_ready();
    }
    tl.stamp();   // 3 layer-1 epilogue done (edge) 
    // L2 prefetch of the next tile's rows (one 128-byte line per thread and stream); its indices
    // were requested a whole layer ago
    {
      const long long ng = grow + (long long)gridDim.x * TILE_ROWS;
      if (ng < p.M) {
        if (!(EDGE && p.rowmap)) prefetch_l2(reinterpret_cast<const char*>(p.in0 + ng * H) + 128 * m.cq);
        if (EDGE) {
          const char* pr = reinterpret_cast<const char*>(
              p.P + (long long)((m.cq & 2) ? nxt.snd_o : nxt.rcv_o) * 256 + ((m.cq & 2) ? 128 : 0));
          prefetch_l2(pr + 128 * (m.cq & 1));
        } else {
          prefetch_l2(reinterpret_cast<const char*>(p.in1 + ng * H) + 128 * m.cq);
        }
      }
    }

    // ---- layers 2, 3
    // node kernel: the hidden activations alternate between the two halves of sA (h1 -> first, h2 -> second, h3 ->
    // first) so that each tile can leave for global memory as one bulk copy (the saves node_bwd bulk-loads) while the
    // next product reads it; thread 0 checks that the copy issued a layer earlier has read its half before it issues
    // the product whose epilogue overwrites that half (every thread then waits for that product anyway)
#pragma unroll 1
    for (int layer = 1; layer < 3; ++layer) {
      uint8_t* act_in = (EDGE || layer == 1) ? sA : sA + TILE_BYTES;
      uint8_t* act_out = (EDGE || layer == 2) ? sA : sA + TILE_BYTES;
      if (m.tid == 0) {
        fence_after_sync();
        if (!EDGE && p.save_h1 != nullptr) {
          bulk_s2g(reinterpret_cast<uint8_t*>(layer == 1 ? p.save_h1 : p.save_h2) + (size_t)tile * TILE_BYTES, act_in,
                   TILE_BYTES, pol_keep);
          bulk_commit();
          if (layer == 2) bulk_wait_read1();
        }
        if (EDGE) umma_tile_ts(tmem_acc, tmem_a, sW_u32 + W1_BYTES + (layer - 1) * TILE_BYTES, 128, 0, 0, 128, 8, 0);
        else umma_tile(tmem_acc, smem_u32(act_in), 128, 0, sW_u32 + W1_BYTES + (layer - 1) * TILE_BYTES, 128, 0, 0, 128, 8, 0);
        mma_commit(&mbar);
      }
      float4 bb[8];
      load_vec32(layer == 1 ? p.b2 : p.b3, m.c0, bb);
      mma_wait();
      tl.stamp();   // 4, 6 product done
      __half* save = layer == 1 ? p.save_h2 : p.save_h3;
      if (EDGE) {
        epi_bias_relu_tmem_q(taddr, taddr_a, bb, m, rvalid ? save : nullptr, grow);
        acts_ready();
      } else {
        epi_bias_relu_q(taddr, bb, act_out, m, nullptr, 0);
        tile_ready();
      }
      tl.stamp();   // 5, 7 epilogue done
    }

    // ---- layer 4 + RMSNorm + residual (+ segmented sum over receivers)
    if (m.tid == 0) {
      fence_after_sync();
      if (!EDGE && p.save_h1 != nullptr) {      // h3 sits in the first half of sA
        bulk_s2g(reinterpret_cast<uint8_t*>(p.save_h3) + (size_t)tile * TILE_BYTES, sA, TILE_BYTES, pol_keep);
        bulk_commit();
      }
      if (EDGE) umma_tile_ts(tmem_acc, tmem_a, sW_u32 + W1_BYTES + 2 * TILE_BYTES, 128, 0, 0, 128, 8, 0);
      else umma_tile(tmem_acc, sA_u32, 128, 0, sW_u32 + W1_BYTES + 2 * TILE_BYTES, 128, 0, 0, 128, 8, 0);
      mma_commit(&mbar);
    }
    float4 old[8];
    int dst[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {       // residual rows (coalesced mapping), in flight during the product
      dst[i] = __shfl_sync(0xffffffffu, dst_c, i);
      old[i] = dst[i] >= 0 ? ld_f4_hint(p.in0 + (long long)dst[i] * H + 4 * m.lane, pol_stream)
                           : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    mma_wait();
    // node kernel: the output staging below reuses both halves of sA -- the copies of h2 and h3 must have read them
    // (the barriers inside row_sum4 publish this to the other threads)
    if (!EDGE && m.tid == 0 && p.save_h1 != nullptr) bulk_wait_read0();
    tl.stamp();   // 8 layer-4 product done (residual rows requested before)
    {
      float v[32];
      tmem_ld32(taddr, v);
      float4 bb[8];
      load_vec32(p.b4, m.c0, bb);
      tmem_ld_wait();
      float ss = 0.f;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        v[4 * j] += bb[j].x; v[4 * j + 1] += bb[j].y; v[4 * j + 2] += bb[j].z; v[4 * j + 3] += bb[j].w;
        ss += v[4 * j] * v[4 * j] + v[4 * j + 1] * v[4 * j + 1] + v[4 * j + 2] * v[4 * j + 2] + v[4 * j + 3] * v[4 * j + 3];
      }
      // (the barriers inside row_sum4 also order the product's reads of sA before the node kernel's staging writes)
      float4 sc[8];
      load_vec32(p.scale, m.c0, sc);
      ss = row_sum4(ss, tail->red, m);
      float inv = 1.0f / (sqrtf(ss) * p.inv_sqrt_d + RMS_EPS);
      if (!EDGE && p.no_norm) {
        inv = 1.0f;
#pragma unroll
        for (int j = 0; j < 8; ++j) sc[j] = make_float4(1.f, 1.f, 1.f, 1.f);
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float4 o;
        o.x = sc[j].x * (v[4 * j + 0] * inv);
        o.y = sc[j].y * (v[4 * j + 1] * inv);
        o.z = sc[j].z * (v[4 * j + 2] * inv);
        o.w = sc[j].w * (v[4 * j + 3] * inv);
        *reinterpret_cast<float4*>(stage + stage_off(m.row, m.c0 + 4 * j)) = o;
      }
    }
    fence_before_sync();
    __syncthreads();
    tl.stamp();   // 9 norm + staging done
    {
      int cur = -1;
      bool first_run = false;
      float4 sum = make_float4(0.f, 0.f, 0.f, 0.f);
      const int prev_c = cur_idx_prev;
      // a run of equal receivers inside this warp's 8 rows: the run that continues from an earlier 8-row slot leaves
      // its partial in carry[slot], every other run stores its sum to its node's row of aggr (aggr_fixup_kernel adds
      // the carries in slot order: deterministic, no atomics)
      auto flush_run = [&](int node, const float4& v, bool first) {
        if (first && prev_c == node) {
          st_f4_hint(p.carry + (tile * 16 + m.warp) * H + 4 * m.lane, v, pol_keep);
        } else {
          st_f4_hint(p.aggr + (long long)node * H + 4 * m.lane, v, pol_keep);
        }
      };
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float4 y = *reinterpret_cast<const float4*>(stage + stage_off(m.warp * 8 + i, 4 * m.lane));
        const long long orow = (EDGE && p.out_csr) ? row0 + m.warp * 8 + i : (long long)dst[i];
        const int rc = EDGE ? __shfl_sync(0xffffffffu, rcv_c, i) : -1;
        if (dst[i] < 0 && EDGE && p.e16_next != nullptr)      // rows past the end: zeros in the image
          *reinterpret_cast<uint2*>(reinterpret_cast<uint8_t*>(p.e16_next) + (size_t)tile * TILE_BYTES +
                                    tile_off(128, m.warp * 8 + i, 4 * m.lane)) = make_uint2(0u, 0u);
        if (dst[i] >= 0) {
          const float4 o = make_float4(old[i].x + y.x, old[i].y + y.y, old[i].z + y.z, old[i].w + y.w);
          if (p.out != nullptr) st_f4_hint(p.out + orow * H + 4 * m.lane, o, pol_stream);
          if (EDGE && p.e16_next != nullptr) {   // the same row in the next step's operand-tile image (8 bytes per lane)
            uint2 pk;
            pk.x = pack_h2(o.x, o.y);
            pk.y = pack_h2(o.z, o.w);
            const size_t ioff = (size_t)tile * TILE_BYTES + tile_off(128, m.warp * 8 + i, 4 * m.lane);
            *reinterpret_cast<uint2*>(reinterpret_cast<uint8_t*>(p.e16_next) + ioff) = pk;
            if (p.e16lo_next != nullptr) {       // what the image rounds away, again in fp16
              const float2 h0 = __half22float2(*reinterpret_cast<const __half2*>(&pk.x));
              const float2 h1 = __half22float2(*reinterpret_cast<const __half2*>(&pk.y));
              uint2 lo;
              lo.x = pack_h2(o.x - h0.x, o.y - h0.y);
              lo.y = pack_h2(o.z - h1.x, o.w - h1.y);
              *reinterpret_cast<uint2*>(reinterpret_cast<uint8_t*>(p.e16lo_next) + ioff) = lo;
            }
          }
          if (EDGE) {
            if (rc != cur) {
              if (cur >= 0) flush_run(cur, sum, first_run);
              first_run = (i == 0);
              cur = rc;
              sum = y;
            } else {
              sum.x += y.x; sum.y += y.y; sum.z += y.z; sum.w += y.w;
            }
          }
        }
      }
      if (EDGE && cur >= 0) flush_run(cur, sum, first_run);
    }
    __syncthreads();      // staging / operand tile free for the next tile
    if (!EDGE && p.ready_out != nullptr && m.tid == 0) tile_publish(p.ready_out, tile);
    tl.stamp();   // 10 output done
  }
  if (!w_ready) mbar_wait(&wbar, 0);   // never exit with the weight copy in flight
  if (!EDGE && m.tid == 0) bulk_wait0();

  fence_before_sync();
  __syncthreads();
  if (m.warp == 0) tmem_dealloc<TMEM_COLS>(tmem_slot);
}

// ---------------------------------------------------------------------------------
// p kernel: P[N,256] (fp16) = x16 * [W_i | W_j]^T + [b1 | 0]   (one N=256 product per tile)
struct PFwdParams {
  long long N;
  const float* x;        // [N,128]
  const __half* Wcat;    // panel tile, 256 rows (out) x 128 cols (in) = 64 KB
  const float* b1;       // [128] edge-block layer-1 bias, folded into P_i
  __half* P;             // [N,256]
  int32_t* ready_in;     // or null: tile-level chaining with the node update launched before this one
  int32_t* sched;        // (mlp_common.cuh): ready flags of x's tiles, {next tile, CTAs done}
};

__global__ void __launch_bounds__(NT, 1) p_fwd_kernel(const PFwdParams p) {
  // The P rows leave through a shared-memory staging tile and one 512-byte bulk copy per row: a row-owner store of the
  // accumulator mapping (a thread owns 32 columns of ONE row) touches 32 distinct lines per instruction and cost more
  // than the product.  Staging rows are padded to 528 bytes so that the 16-byte chunks of consecutive rows fall into
  // different banks (conflict-free row-owner writes); the copies are issued by threads 0..127, one row each.
  constexpr uint32_t PITCH = 528;
  extern __shared__ __align__(1024) uint8_t smem[];
  check_smem_alignment(smem);
  uint8_t* sW = smem;                   // 64 KB
  uint8_t* sA = smem + 2 * TILE_BYTES;  // 32 KB
  uint8_t* sP = smem + 3 * TILE_BYTES;  // 128 x 528 B staging of the output rows
  Tail2* tail = reinterpret_cast<Tail2*>(smem + 3 * TILE_BYTES + 128 * PITCH);
  uint64_t& wbar = tail->bar[0];
  uint64_t& mbar = tail->bar[1];
  uint32_t& tmem_slot = tail->tmem_slot;
  const Map m = make_map();
  if (m.tid == 0) {
    mbar_init(&wbar, 1);
    mbar_init(&mbar, 1);
    mbar_init_fence();
  }
  if (m.warp == 0) tmem_alloc<256>(&tmem_slot);
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const bool chained = p.ready_in != nullptr;
  if (!chained) pdl_wait();   // the previous kernel in the stream is complete: global memory may be touched
  pdl_launch();               // (common.cuh) the next kernel may start its own prologue under this one
  const uint32_t tmem_acc = tmem_slot;
  if (m.tid == 0) {
    mbar_arrive_expect_tx(&wbar, 2 * TILE_BYTES);
    bulk_g2s(sW, p.Wcat, TILE_BYTES, &wbar);
    bulk_g2s(sW + TILE_BYTES, reinterpret_cast<const uint8_t*>(p.Wcat) + TILE_BYTES, TILE_BYTES, &wbar);
  }
  float4 bb[8];
  load_vec32(p.b1, m.c0, bb);
  const uint64_t pol_keep = policy_evict_last();
  bool w_ready = false;
  uint32_t phase = 0;
  const long long num_tiles = (p.N + TILE_ROWS - 1) / TILE_ROWS;
  int* s_tile = reinterpret_cast<int*>(tail->red);
  for (long long tile = blockIdx.x;; tile += gridDim.x) {
    if (chained) {      // next tile in index order = the order the node update finishes them in; wait for its rows
      if (m.tid == 0) {
        const int t = atomicAdd(p.sched, 1);
        if (t < num_tiles) tile_acquire(p.ready_in, t);
        *s_tile = t;
      }
      __syncthreads();
      tile = *s_tile;   // (thread 0 writes it again only after the barriers below)
    }
    if (tile >= num_tiles) break;
    const long long row0 = tile * TILE_ROWS;
    auto rfx = [&](int i) -> long long {
      const long long g = row0 + m.warp * 8 + i;
      return g < p.N ? g : -1;
    };
    if (chained) load8_f32_to_tile_cg(sA, 128, 0, p.x, m, rfx);
    else load8_f32_to_tile(sA, 128, 0, p.x, m, rfx, pol_keep);
    if (!chained && m.tid == 0 && tile + gridDim.x < num_tiles) {      // next tile's rows into L2: one bulk prefetch
      const long long n0 = (tile + gridDim.x) * TILE_ROWS;
      const long long nr = (p.N - n0) < TILE_ROWS ? (p.N - n0) : TILE_ROWS;
      bulk_prefetch_l2(p.x + n0 * H, (uint32_t)(nr * 512));
    }
    fence_async_smem();
    fence_before_sync();
    __syncthreads();
    if (!w_ready) {
      mbar_wait(&wbar, 0);
      w_ready = true;
    }
    if (m.tid == 0) {
      fence_after_sync();
      umma_tile(tmem_acc, smem_u32(sA), 128, 0, smem_u32(sW), 256, 0, 0, 256, 8, 0);
      mma_commit(&mbar);
    }
    if (m.tid < 128) bulk_wait_read0();      // this thread's row copy of the previous tile has left the staging tile
    mbar_wait(&mbar, phase);
    phase ^= 1;
    fence_after_sync();
    __syncthreads();                         // (publishes the wait above to the threads that write the staging rows)
    const long long grow = row0 + m.row;
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      float v[32];
      tmem_ld32(tmem_acc + m.lane_off + half * 128 + m.c0, v);
      tmem_ld_wait();
      if (half == 0) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          v[4 * j] += bb[j].x; v[4 * j + 1] += bb[j].y; v[4 * j + 2] += bb[j].z; v[4 * j + 3] += bb[j].w;
        }
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        uint4 pk;
        pk.x = pack_h2(v[8 * j + 0], v[8 * j + 1]);
        pk.y = pack_h2(v[8 * j + 2], v[8 * j + 3]);
        pk.z = pack_h2(v[8 * j + 4], v[8 * j + 5]);
        pk.w = pack_h2(v[8 * j + 6], v[8 * j + 7]);
        *reinterpret_cast<uint4*>(sP + m.row * PITCH + (half * 128 + m.c0 + 8 * j) * 2) = pk;
      }
    }
    fence_async_smem();
    fence_before_sync();
    __syncthreads();
    if (m.tid < 128 && row0 + m.tid < p.N) {
      bulk_s2g(p.P + (row0 + m.tid) * 256, sP + m.tid * PITCH, 512, pol_keep);
      bulk_commit();
    }
    (void)grow;
  }
  if (chained) {
    if (m.tid == 0) sched_done(p.sched);
    pdl_wait();       // the node update is complete: what this launch's successors rely on
  }
  if (!w_ready) mbar_wait(&wbar, 0);
  if (m.tid < 128) bulk_wait0();
  fence_before_sync();
  __syncthreads();
  if (m.warp == 0) tmem_dealloc<256>(tmem_slot);
}

// ---------------------------------------------------------------------------------
// weight packing: fp32 nn.Linear weights [out,in] -> fp16 panel tiles (mlp_fwd_abi.cuh)
#include "mlp_fwd_abi.cuh"

struct PackParams {
  const float* eW[4];   // edge_block.{0,2,4,6}.weight: [128,384], [128,128] x3
  const float* nW[4];   // node_block.{0,2,4,6}.weight: [128,256], [128,128] x3
  uint8_t* out;
};

__global__ void pack_block_kernel(const PackParams p) {
  pdl_wait();
  pdl_launch();
  // one thread per 8 consecutive columns (one 16-byte chunk)
  const int total_chunks = (int)(PK_BYTES / 16);
  for (int c = blockIdx.x * blockDim.x + threadIdx.x; c < total_chunks; c += gridDim.x * blockDim.x) {
    const int64_t byte = (int64_t)c * 16;
    const float* src;
    int ld, rows, cols, coff = 0, roff = 0;
    int64_t base;
    if (byte < PK_WCAT) {
      int ti = (int)(byte / 32768);
      base = (int64_t)ti * 32768; rows = 128; cols = 128;
      src = p.eW[ti]; ld = ti == 0 ? 384 : 128;
    } else if (byte < PK_NODE) {
      base = PK_WCAT; rows = 256; cols = 128; src = p.eW[0]; ld = 384;
    } else if (byte < PK_NODE + 65536) {
      base = PK_NODE; rows = 128; cols = 256; src = p.nW[0]; ld = 256;
    } else {
      int ti = (int)((byte - PK_NODE - 65536) / 32768);
      base = PK_NODE + 65536 + (int64_t)ti * 32768; rows = 128; cols = 128; src = p.nW[1 + ti]; ld = 128;
    }
    // invert tile_off: byte offset inside tile -> (row, col of first element of the chunk)
    const uint32_t off = (uint32_t)(byte - base);
    const uint32_t panel = off / (rows * 128), rem = off % (rows * 128);
    const uint32_t row = rem / 128, pchunk = (rem % 128) / 16;
    const uint32_t col = panel * 64 + ((pchunk ^ (row & 7)) << 3);
    int srow = row, scol = col;
    if (base == PK_WCAT) {            // rows 0..127: W1[:,128:256]; rows 128..255: W1[:,256:384]
      srow = row & 127;
      scol = col + (row < 128 ? 128 : 256);
    }
    (void)cols; (void)coff; (void)roff;
    const float* s = src + (int64_t)srow * ld + scol;
    float4 a = __ldg(reinterpret_cast<const float4*>(s)), b = __ldg(reinterpret_cast<const float4*>(s + 4));
    uint4 pk;
    pk.x = pack_h2(a.x, a.y); pk.y = pack_h2(a.z, a.w); pk.z = pack_h2(b.x, b.y); pk.w = pack_h2(b.z, b.w);
    *reinterpret_cast<uint4*>(p.out + byte) = pk;
  }
}

// ---- all blocks of a processor in one launch: grid.y = block index
static constexpr int PACK_MAX_BLOCKS = 32;
struct PackManyParams {
  const float* w[PACK_MAX_BLOCKS][8];   // per block: edge_block.{0,2,4,6}.weight, node_block.{0,2,4,6}.weight
  uint8_t* out;                         // [K][PK_BYTES]
};

__global__ void pack_blocks_kernel(const PackManyParams q) {
  pdl_wait();
  pdl_launch();
  const int k = blockIdx.y;
  PackParams p;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    p.eW[i] = q.w[k][i];
    p.nW[i] = q.w[k][4 + i];
  }
  p.out = q.out + (size_t)k * PK_BYTES;
  const int total_chunks = (int)(PK_BYTES / 16);
  for (int c = blockIdx.x * blockDim.x + threadIdx.x; c < total_chunks; c += gridDim.x * blockDim.x) {
    const int64_t byte = (int64_t)c * 16;
    const float* src;
    int ld, rows;
    int64_t base;
    if (byte < PK_WCAT) {
      int ti = (int)(byte / 32768);
      base = (int64_t)ti * 32768; rows = 128;
      src = p.eW[ti]; ld = ti == 0 ? 384 : 128;
    } else if (byte < PK_NODE) {
      base = PK_WCAT; rows = 256; src = p.eW[0]; ld = 384;
    } else if (byte < PK_NODE + 65536) {
      base = PK_NODE; rows = 128; src = p.nW[0]; ld = 256;
    } else {
      int ti = (int)((byte - PK_NODE - 65536) / 32768);
      base = PK_NODE + 65536 + (int64_t)ti * 32768; rows = 128; src = p.nW[1 + ti]; ld = 128;
    }
    const uint32_t off = (uint32_t)(byte - base);
    const uint32_t panel = off / (rows * 128), rem = off % (rows * 128);
    const uint32_t row = rem / 128, pchunk = (rem % 128) / 16;
    const uint32_t col = panel * 64 + ((pchunk ^ (row & 7)) << 3);
    int srow = row, scol = col;
    if (base == PK_WCAT) {
      srow = row & 127;
      scol = col + (row < 128 ? 128 : 256);
    }
    const float* sp = src + (int64_t)srow * ld + scol;
    float4 a = __ldg(reinterpret_cast<const float4*>(sp)), b = __ldg(reinterpret_cast<const float4*>(sp + 4));
    uint4 pk;
    pk.x = pack_h2(a.x, a.y); pk.y = pack_h2(a.z, a.w); pk.z = pack_h2(b.x, b.y); pk.w = pack_h2(b.z, b.w);
    *reinterpret_cast<uint4*>(p.out + byte) = pk;
  }
}

// K blocks in one launch per 32 blocks.  weights_host: HOST array of K * 8 DEVICE pointers, per block
// edge_block.{0,2,4,6}.weight then node_block.{0,2,4,6}.weight (fp32, the nn.Linear layout); packed: [K][packed bytes].
extern "C" int mgn_pack_blocks(int K, const float* const* weights_host, void* packed, void* stream) {
  MGN_CHECK_ARG(K > 0 && weights_host && packed, "pack_blocks: bad args");
  for (int k0 = 0; k0 < K; k0 += PACK_MAX_BLOCKS) {
    const int kn = std::min(PACK_MAX_BLOCKS, K - k0);
    PackManyParams q{};
    for (int k = 0; k < kn; ++k)
      for (int i = 0; i < 8; ++i) {
        q.w[k][i] = weights_host[(size_t)(k0 + k) * 8 + i];
        MGN_CHECK_ARG(q.w[k][i] != nullptr, "pack_blocks: null weight pointer");
      }
    q.out = (uint8_t*)packed + (size_t)k0 * PK_BYTES;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(22, (unsigned)kn);
    cfg.blockDim = dim3(256);
    cfg.stream = (cudaStream_t)stream;
    MGN_CUDA(cudaLaunchKernelEx(&cfg, pack_blocks_kernel, q));
    MGN_LAUNCH_CHECK();
  }
  return MGN_OK;
}

extern "C" int64_t mgn_packed_block_bytes(void) { return PK_BYTES; }

extern "C" int mgn_pack_block(const float* eW1, const float* eW2, const float* eW3, const float* eW4,
                              const float* nW1, const float* nW2, const float* nW3, const float* nW4,
                              void* packed, void* stream) {
  MGN_CHECK_ARG(eW1 && eW2 && eW3 && eW4 && nW1 && nW2 && nW3 && nW4 && packed, "pack_block: null pointer");
  PackParams p{{eW1, eW2, eW3, eW4}, {nW1, nW2, nW3, nW4}, (uint8_t*)packed};
  MGN_CUDA(mgn_launch(pack_block_kernel, 88, 256, 0, (cudaStream_t)stream, p));
  MGN_LAUNCH_CHECK();
  return MGN_OK;
}

static constexpr int EDGE_FWD_SMEM = 4 * 32768 + 32768 + 2 * 32768 + TAIL_BYTES;
static constexpr int NODE_FWD_SMEM = 5 * 32768 + 2 * 32768 + TAIL_BYTES;
static constexpr int P_FWD_SMEM = 3 * 32768 + 128 * 528 + TAIL_BYTES;

// GraphNetBlock edge update + aggregation (layers.py:687-698).  e_in/e_out [E,128] fp32 in
// the caller's edge order when rowmap (= CSR perm) is given, else already in CSR order.
// carry ([ceil(E/8),128] fp32 workspace or NULL) and norm_dim: as in mgn_edge_fwd_hl (edge_fwd_hl.cu).
// mgn_edge_fwd2 = mgn_edge_fwd plus e16lo_next (tile images of e_out - fp16(e_out), the second half of the image +
// remainder form of the residual stream that mgn_edge_fwd_hl consumes); with both images requested e_out may be NULL.
extern "C" int mgn_edge_fwd2(int64_t E, int64_t N, const float* e_in, const int32_t* rowmap, float* e_out,
                             const void* P, const int32_t* senders, const int32_t* receivers, float* aggr,
                             const void* packed, const float* b2, const float* b3, const float* b4,
                             const float* scale, void* save_h1, void* save_h2, void* save_h3, void* e16_out,
                             const void* e16_in, void* e16_next, void* e16lo_next, int out_in_csr_order,
                             float* carry, const int32_t* rowptr, int norm_dim, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  MGN_CHECK_ARG(E >= 0 && N > 0 && aggr, "edge_fwd: bad sizes");
  if (rowptr == nullptr || E == 0) MGN_CUDA(cudaMemsetAsync(aggr, 0, sizeof(float) * N * H, stream));
  if (E == 0) return MGN_OK;
  MGN_CHECK_ARG(e_in && P && senders && receivers && packed && b2 && b3 && b4 && scale && carry,
                "edge_fwd: null pointer");
  MGN_CHECK_ARG(e_out || (e16_next && e16lo_next), "edge_fwd: e_out may be null only if both output images are given");
  MGN_CHECK_ARG(!e16lo_next || e16_next, "edge_fwd: e16lo_next needs e16_next");
  MGN_CHECK_ARG(e_in != e_out, "edge_fwd: in-place not supported");
  FwdParams p{};
  p.M = E; p.in0 = e_in; p.rowmap = rowmap; p.out_csr = (rowmap != nullptr && out_in_csr_order) ? 1 : 0;
  p.out = e_out; p.P = (const __half*)P;
  p.senders = senders; p.receivers = receivers; p.aggr = aggr;
  p.W = reinterpret_cast<const __half*>(reinterpret_cast<const uint8_t*>(packed) + PK_EDGE);
  p.b2 = b2; p.b3 = b3; p.b4 = b4; p.scale = scale;
  p.save_h1 = (__half*)save_h1; p.save_h2 = (__half*)save_h2; p.save_h3 = (__half*)save_h3;
  p.e16_out = (__half*)e16_out; p.e16_in = (const __half*)e16_in; p.e16_next = (__half*)e16_next;
  p.e16lo_next = (__half*)e16lo_next;
  MGN_CHECK_ARG(norm_dim > 0 && norm_dim <= H, "edge_fwd: norm_dim must be in 1..128");
  p.carry = carry; p.inv_sqrt_d = 1.0f / sqrtf((float)norm_dim);
  p.dbg = g_mgn_timeline_host; p.dbg_cap = g_mgn_timeline_cap_host;
  static bool attr = false;
  if (!attr) {
    MGN_CUDA(cudaFuncSetAttribute(mlp_fwd_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, EDGE_FWD_SMEM));
    attr = true;
  }
  const int64_t tiles = ceil_div64(E, TILE_ROWS);
  const int grid = (int)std::min<int64_t>(mgn_num_sms(), tiles);
  MGN_CUDA(mgn_launch(mlp_fwd_kernel<true>, grid, NT, EDGE_FWD_SMEM, stream, p));
  MGN_LAUNCH_CHECK();
  return mgn_launch_aggr_fixup(E, N, 8, receivers, rowptr, carry, aggr, stream);
}

// GraphNetBlock node update (layers.py:733-746, 699): x_out = x + node_block(cat[x, aggr]).
// flags bit 0: no RMSNorm (x_out = x + MLP(...) with the last layer's bias only; `scale` is ignored but must be valid)
extern "C" int mgn_node_fwd3(int64_t N, const float* x, const float* aggr, float* x_out, const void* packed,
                             const float* b1, const float* b2, const float* b3, const float* b4,
                             const float* scale, void* save_h1, void* save_h2, void* save_h3, int flags,
                             int norm_dim, int32_t* ready_out, void* stream_);
extern "C" int mgn_node_fwd2(int64_t N, const float* x, const float* aggr, float* x_out, const void* packed,
                             const float* b1, const float* b2, const float* b3, const float* b4,
                             const float* scale, void* save_h1, void* save_h2, void* save_h3, int flags,
                             int norm_dim, void* stream_) {
  return mgn_node_fwd3(N, x, aggr, x_out, packed, b1, b2, b3, b4, scale, save_h1, save_h2, save_h3, flags, norm_dim,
                       nullptr, stream_);
}
extern "C" int mgn_node_fwd3(int64_t N, const float* x, const float* aggr, float* x_out, const void* packed,
                             const float* b1, const float* b2, const float* b3, const float* b4,
                             const float* scale, void* save_h1, void* save_h2, void* save_h3, int flags,
                             int norm_dim, int32_t* ready_out, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  MGN_CHECK_ARG(N > 0 && x && aggr && x_out && packed && b1 && b2 && b3 && b4 && scale, "node_fwd: bad args");
  MGN_CHECK_ARG(x != x_out, "node_fwd: in-place not supported");
  MGN_CHECK_ARG((save_h1 == nullptr) == (save_h2 == nullptr) && (save_h1 == nullptr) == (save_h3 == nullptr),
                "node_fwd: the three saves go together");
  FwdParams p{};
  p.M = N; p.in0 = x; p.in1 = aggr; p.out = x_out;
  p.W = reinterpret_cast<const __half*>(reinterpret_cast<const uint8_t*>(packed) + PK_NODE);
  p.b1 = b1; p.b2 = b2; p.b3 = b3; p.b4 = b4; p.scale = scale;
  p.save_h1 = (__half*)save_h1; p.save_h2 = (__half*)save_h2; p.save_h3 = (__half*)save_h3;
  p.no_norm = flags & 1;
  p.ready_out = ready_out;
  MGN_CHECK_ARG(norm_dim > 0 && norm_dim <= H, "node_fwd: norm_dim must be in 1..128");
  p.inv_sqrt_d = 1.0f / sqrtf((float)norm_dim);
  p.dbg = g_mgn_timeline_host; p.dbg_cap = g_mgn_timeline_cap_host;
  static bool attr = false;
  if (!attr) {
    MGN_CUDA(cudaFuncSetAttribute(mlp_fwd_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, NODE_FWD_SMEM));
    attr = true;
  }
  const int grid = (int)std::min<int64_t>(mgn_num_sms(), ceil_div64(N, TILE_ROWS));
  MGN_CUDA(mgn_launch(mlp_fwd_kernel<false>, grid, NT, NODE_FWD_SMEM, stream, p));
  MGN_LAUNCH_CHECK();
  return MGN_OK;
}

extern "C" int mgn_node_fwd(int64_t N, const float* x, const float* aggr, float* x_out, const void* packed,
                            const float* b1, const float* b2, const float* b3, const float* b4,
                            const float* scale, void* save_h1, void* save_h2, void* save_h3, void* stream_) {
  return mgn_node_fwd2(N, x, aggr, x_out, packed, b1, b2, b3, b4, scale, save_h1, save_h2, save_h3, 0, H, stream_);
}

// per-node layer-1 partial products of the edge block: P = [x W_i^T + b1 | x W_j^T] (fp16)
extern "C" int mgn_p_fwd_c(int64_t N, const float* x, const void* packed, const float* b1_edge, void* P,
                           int32_t* ready_in, int32_t* sched, void* stream_);
extern "C" int mgn_p_fwd(int64_t N, const float* x, const void* packed, const float* b1_edge, void* P,
                         void* stream_) {
  return mgn_p_fwd_c(N, x, packed, b1_edge, P, nullptr, nullptr, stream_);
}
extern "C" int mgn_p_fwd_c(int64_t N, const float* x, const void* packed, const float* b1_edge, void* P,
                           int32_t* ready_in, int32_t* sched, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  MGN_CHECK_ARG(N > 0 && x && packed && b1_edge && P, "p_fwd: bad args");
  MGN_CHECK_ARG((ready_in == nullptr) == (sched == nullptr), "p_fwd: ready_in and sched go together");
  PFwdParams p{N, x, reinterpret_cast<const __half*>(reinterpret_cast<const uint8_t*>(packed) + PK_WCAT),
               b1_edge, (__half*)P, ready_in, sched};
  static bool attr = false;
  if (!attr) {
    MGN_CUDA(cudaFuncSetAttribute(p_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, P_FWD_SMEM));
    attr = true;
  }
  const int grid = (int)std::min<int64_t>(mgn_num_sms(), ceil_div64(N, TILE_ROWS));
  MGN_CUDA(mgn_launch(p_fwd_kernel, grid, NT, P_FWD_SMEM, stream, p));
  MGN_LAUNCH_CHECK();
  return MGN_OK;
}
