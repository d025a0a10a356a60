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
// summed in the registers of a second warpgroup.  Weight tiles stream through two 32 KB shared
// memory slots (bulk async copies from L2), the remaining five slots hold the activation /
// gradient tiles of the current 128-row tile.  Gradients arrive pre-scaled by a power of two
// (see processor.py) so that fp16 operands keep their precision; everything here is linear in g.
#include "mlp_common.cuh"
#include "mlp_fwd_abi.cuh"

using namespace mgn;

static constexpr uint32_t SLOT = 32768;

__device__ __forceinline__ void load_w_slot(uint8_t* slot, const __half* W, int widx, uint64_t* bar) {
  mbar_arrive_expect_tx(bar, SLOT);
  bulk_g2s(slot, reinterpret_cast<const uint8_t*>(W) + (size_t)widx * SLOT, SLOT, bar);
}

// ===================================================================================== edge
struct EdgeBwdParams {
  long long E;
  const float* e_in;        // [E,128] step input (saved)
  const int32_t* rowmap;    // CSR position -> row of e_in / g_eout / g_e (null = identity)
  const float* g_eout;      // [E,128] or null (= 0)
  const float* g_aggr;      // [N,128]
  const __half* P;          // [N,256]
  const int32_t* senders;
  const int32_t* receivers;
  float* g_e;               // [E,128] out
  float* gP;                // [N,256] fp32, zero on entry
  const __half* W;          // packed edge tiles We, W2, W3, W4
  const float* b2;
  const float* b3;
  const float* b4;
  const float* scale;
  float* grad;              // flat gradient block (GB_* offsets), accumulated with red.add
};

__global__ void __launch_bounds__(256, 1) edge_bwd_kernel(const EdgeBwdParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  check_smem_alignment(smem);
  uint8_t* sWA = smem;
  uint8_t* sWB = smem + SLOT;
  uint8_t* sE = smem + 2 * SLOT;
  uint8_t* sH1 = smem + 3 * SLOT;
  uint8_t* sH2 = smem + 4 * SLOT;
  uint8_t* sH3 = smem + 5 * SLOT;
  uint8_t* sG = smem + 6 * SLOT;
  uint8_t* stage = sH3;   // 64 KB fp32 staging = sH3 + sG (both dead by then)
  SmemTail* tail = reinterpret_cast<SmemTail*>(smem + 7 * SLOT);
  uint64_t& bar_wA = tail->bar[0];
  uint64_t& bar_wB = tail->bar[1];
  uint64_t& bar_mma = tail->bar[2];
  uint64_t& bar_dwe = tail->bar[3];
  uint64_t& bar_wfree = tail->bar[4];
  uint32_t& tmem_slot = tail->tmem_slot;
  const float* sb2 = tail->vec[1];
  const float* sb3 = tail->vec[2];
  const float* sb4 = tail->vec[3];
  const float* ssc = tail->vec[4];

  const int tid = threadIdx.x;
  const int wg = tid >> 7;
  const int t = tid & 127;
  const int w = t >> 5;
  const int lane = tid & 31;

  copy_vec128(tail->vec[1], p.b2, tid, blockDim.x);
  copy_vec128(tail->vec[2], p.b3, tid, blockDim.x);
  copy_vec128(tail->vec[3], p.b4, tid, blockDim.x);
  copy_vec128(tail->vec[4], p.scale, tid, blockDim.x);
  if (tid == 0) {
    mbar_init(&bar_wA, 1);
    mbar_init(&bar_wB, 1);
    mbar_init(&bar_mma, 1);
    mbar_init(&bar_dwe, 1);
    mbar_init(&bar_wfree, 128);
    mbar_init_fence();
  }
  if ((tid >> 5) == 0) tmem_alloc<512>(&tmem_slot);
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t lane_off = (uint32_t)(w * 32) << 16;
  const uint32_t tm_W = tmem_slot, tm_dW4 = tmem_slot + 128, tm_dW3 = tmem_slot + 256, tm_dW2 = tmem_slot + 384;
  const long long num_tiles = (p.E + TILE_ROWS - 1) / TILE_ROWS;

  if (wg == 1) {
    // ---------------- register accumulator of dWe = sum over tiles of dz1^T e16
    float acc[128];
#pragma unroll
    for (int j = 0; j < 128; ++j) acc[j] = 0.f;
    uint32_t it = 0;
    for (long long tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
      mbar_wait(&bar_dwe, it & 1);
      fence_after_sync();
#pragma unroll
      for (int cq = 0; cq < 4; ++cq) {
        float v[32];
        tmem_ld32(tm_W + lane_off + cq * 32, v);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j) acc[cq * 32 + j] += v[j];
      }
      fence_before_sync();
      mbar_arrive(&bar_wfree);
    }
    float* dst = p.grad + GB_E_W1 + (long long)t * 384;
#pragma unroll
    for (int j = 0; j < 32; ++j)
      red_add_f4(dst + 4 * j, make_float4(acc[4 * j], acc[4 * j + 1], acc[4 * j + 2], acc[4 * j + 3]));
  } else {
    // ---------------- compute warpgroup
    const uint32_t uWA = smem_u32(sWA), uWB = smem_u32(sWB), uE = smem_u32(sE), uH1 = smem_u32(sH1),
                   uH2 = smem_u32(sH2), uH3 = smem_u32(sH3), uG = smem_u32(sG);
    uint32_t phase = 0;
    float dsc[4] = {0.f, 0.f, 0.f, 0.f};
    float db2a = 0.f, db2b = 0.f, db3a = 0.f, db3b = 0.f, db4a = 0.f, db4b = 0.f;
    if (t == 0) {
      load_w_slot(sWA, p.W, 0, &bar_wA);
      load_w_slot(sWB, p.W, 1, &bar_wB);
    }
    auto mma_wait = [&]() {
      mbar_wait(&bar_mma, phase);
      phase ^= 1;
      fence_after_sync();
    };
    auto tile_ready = [&]() {
      fence_async_smem();
      fence_before_sync();
      named_bar(1, 128);
    };
    const int hw = lane >> 4, q = lane & 15;
    uint32_t it = 0;
    for (long long tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
      const long long row0 = tile * TILE_ROWS;
      const long long grow = row0 + t;
      const bool rvalid = grow < p.E;
      int rcv = -1, snd = -1, my_dst = -1;
      if (rvalid) {
        rcv = __ldg(p.receivers + grow);
        snd = __ldg(p.senders + grow);
        my_dst = p.rowmap ? __ldg(p.rowmap + grow) : (int)grow;
      }
      // next tile's indices (used only for L2 prefetches, consumed after the first product is issued)
      const long long ngrow = grow + (long long)gridDim.x * TILE_ROWS;
      int nrcv = -1, nsnd = -1, ndst = -1;
      if (ngrow < p.E) {
        nrcv = __ldg(p.receivers + ngrow);
        nsnd = __ldg(p.senders + ngrow);
        ndst = p.rowmap ? __ldg(p.rowmap + ngrow) : (int)ngrow;
      }

      // ---- load: e -> sE (fp16), g = g_eout + g_aggr[recv] -> sG (fp16)
      load_rows_f32_to_tile_b<16>(sE, 128, 0, p.e_in, w * 32, 32, lane, [&](int r) -> long long {
        return (long long)__shfl_sync(0xffffffffu, my_dst, r & 31);
      });
      if (p.g_eout) {
        load_rows_sum2_to_tile<8>(sG, 0, p.g_eout, p.g_aggr, w * 32, 32, lane,
                                  [&](int r) -> long long { return (long long)__shfl_sync(0xffffffffu, my_dst, r & 31); },
                                  [&](int r) -> long long { return (long long)__shfl_sync(0xffffffffu, rcv, r & 31); });
      } else {
        load_rows_f32_to_tile_b<16>(sG, 128, 0, p.g_aggr, w * 32, 32, lane, [&](int r) -> long long {
          return (long long)__shfl_sync(0xffffffffu, rcv, r & 31);
        });
      }
      tile_ready();

      // ---- R1: z1 = e We^T + P_i[recv] + P_j[send]
      if (t == 0) {
        if (it == 0) mbar_wait(&bar_wA, 0);
        else mbar_wait(&bar_wfree, (it - 1) & 1);
        fence_after_sync();
        umma_tile(tm_W, uE, 128, 0, uWA, 128, 0, 0, 128, 8, 0);
        mma_commit(&bar_mma);
      }
      {
        uint4 pi[16], pj[16];
        if (rvalid) {
          const __half* Pi = p.P + (long long)rcv * 256;
          const __half* Pj = p.P + (long long)snd * 256 + 128;
#pragma unroll
          for (int j = 0; j < 16; ++j) { pi[j] = ldg_u4(Pi + 8 * j); pj[j] = ldg_u4(Pj + 8 * j); }
        } else {
#pragma unroll
          for (int j = 0; j < 16; ++j) { pi[j] = make_uint4(0, 0, 0, 0); pj[j] = pi[j]; }
        }
        // prefetch the next tile's rows into L2
        if (ndst >= 0) {
          prefetch_row_f32(p.e_in + (long long)ndst * H);
          if (p.g_eout) prefetch_row_f32(p.g_eout + (long long)ndst * H);
          prefetch_row_f32(p.g_aggr + (long long)nrcv * H);
          prefetch_row_f16(p.P + (long long)nrcv * 256);
          prefetch_row_f16(p.P + (long long)nsnd * 256 + 128);
        }
        mma_wait();
        if (t == 0) load_w_slot(sWA, p.W, 2, &bar_wA);   // W3 -> A
#pragma unroll
        for (int cq = 0; cq < 4; ++cq) {
          float v[32];
          tmem_ld32(tm_W + lane_off + cq * 32, v);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            float a[8], b[8];
            unpack8(pi[cq * 4 + j], a);
            unpack8(pj[cq * 4 + j], b);
#pragma unroll
            for (int i = 0; i < 8; ++i) v[8 * j + i] = fmaxf(v[8 * j + i] + a[i] + b[i], 0.f);
          }
          store_row_chunk_f16(sH1, 128, t, cq * 32, v);
        }
      }
      tile_ready();
      // ---- R2
      if (t == 0) {
        if (it == 0) mbar_wait(&bar_wB, 0);
        fence_after_sync();
        umma_tile(tm_W, uH1, 128, 0, uWB, 128, 0, 0, 128, 8, 0);
        mma_commit(&bar_mma);
      }
      mma_wait();
      if (t == 0) load_w_slot(sWB, p.W, 3, &bar_wB);     // W4 -> B
      epi_bias_relu(tm_W + lane_off, sb2, sH2, t);
      tile_ready();
      // ---- R3
      if (t == 0) {
        mbar_wait(&bar_wA, 1);
        fence_after_sync();
        umma_tile(tm_W, uH2, 128, 0, uWA, 128, 0, 0, 128, 8, 0);
        mma_commit(&bar_mma);
      }
      mma_wait();
      epi_bias_relu(tm_W + lane_off, sb3, sH3, t);
      tile_ready();
      // ---- R4 + RMSNorm backward
      if (t == 0) {
        mbar_wait(&bar_wB, 1);
        fence_after_sync();
        umma_tile(tm_W, uH3, 128, 0, uWB, 128, 0, 0, 128, 8, 0);
        mma_commit(&bar_mma);
      }
      mma_wait();
      epi_rmsnorm_bwd(tm_W + lane_off, sb4, ssc, sG, t, lane, dsc);
      tile_ready();
      // ---- D4: g_h3 = dz4 W4 ; dW4 += dz4^T h3
      if (t == 0) {
        fence_after_sync();
        umma_tile(tm_W, uG, 128, 0, uWB, 128, 1, 0, 128, 8, 0);
        umma_tile(tm_dW4, uG, 128, 1, uH3, 128, 1, 0, 128, 8, it > 0);
        mma_commit(&bar_mma);
      }
      tile_colsum_pair(sG, t, db4a, db4b);
      mma_wait();
      if (t == 0) load_w_slot(sWB, p.W, 1, &bar_wB);     // W2 -> B
      epi_relu_mask_inplace(tm_W + lane_off, sH3, t);
      tile_ready();
      // ---- D3
      if (t == 0) {
        fence_after_sync();
        umma_tile(tm_W, uH3, 128, 0, uWA, 128, 1, 0, 128, 8, 0);
        umma_tile(tm_dW3, uH3, 128, 1, uH2, 128, 1, 0, 128, 8, it > 0);
        mma_commit(&bar_mma);
      }
      tile_colsum_pair(sH3, t, db3a, db3b);
      mma_wait();
      if (t == 0) load_w_slot(sWA, p.W, 0, &bar_wA);     // We -> A
      epi_relu_mask_inplace(tm_W + lane_off, sH2, t);
      tile_ready();
      // ---- D2
      if (t == 0) {
        mbar_wait(&bar_wB, 0);
        fence_after_sync();
        umma_tile(tm_W, uH2, 128, 0, uWB, 128, 1, 0, 128, 8, 0);
        umma_tile(tm_dW2, uH2, 128, 1, uH1, 128, 1, 0, 128, 8, it > 0);
        mma_commit(&bar_mma);
      }
      tile_colsum_pair(sH2, t, db2a, db2b);
      mma_wait();
      epi_relu_mask_inplace(tm_W + lane_off, sH1, t);    // sH1 <- dz1
      tile_ready();
      // ---- D1: g_e = g_eout + dz1 We ; scatter dz1 while the product runs
      if (t == 0) {
        mbar_wait(&bar_wA, 0);
        fence_after_sync();
        umma_tile(tm_W, uH1, 128, 0, uWA, 128, 1, 0, 128, 8, 0);
        mma_commit(&bar_mma);
      }
      {
        const int cbase = w * 32 + hw * 16;
        int cur = -1;
        float sum[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) sum[i] = 0.f;
#pragma unroll 4
        for (int i = 0; i < 16; ++i) {
          const int r = cbase + i;
          const int rc = __shfl_sync(0xffffffffu, rcv, hw * 16 + i);
          const int sn = __shfl_sync(0xffffffffu, snd, hw * 16 + i);
          const uint4 u = *reinterpret_cast<const uint4*>(sH1 + tile_off(128, r, 8 * q));
          float f[8];
          unpack8(u, f);
          if (rc >= 0) {
            float* dj = p.gP + (long long)sn * 256 + 128 + 8 * q;
            red_add_f4(dj, make_float4(f[0], f[1], f[2], f[3]));
            red_add_f4(dj + 4, make_float4(f[4], f[5], f[6], f[7]));
            if (rc != cur) {
              if (cur >= 0) {
                float* di = p.gP + (long long)cur * 256 + 8 * q;
                red_add_f4(di, make_float4(sum[0], sum[1], sum[2], sum[3]));
                red_add_f4(di + 4, make_float4(sum[4], sum[5], sum[6], sum[7]));
              }
              cur = rc;
#pragma unroll
              for (int k = 0; k < 8; ++k) sum[k] = f[k];
            } else {
#pragma unroll
              for (int k = 0; k < 8; ++k) sum[k] += f[k];
            }
          }
        }
        if (cur >= 0) {
          float* di = p.gP + (long long)cur * 256 + 8 * q;
          red_add_f4(di, make_float4(sum[0], sum[1], sum[2], sum[3]));
          red_add_f4(di + 4, make_float4(sum[4], sum[5], sum[6], sum[7]));
        }
      }
      mma_wait();
      acc_to_stage(tm_W + lane_off, stage, t);
      fence_before_sync();
      named_bar(1, 128);
      // dWe tile product into the (now free) working accumulator; summed by warpgroup 1
      if (t == 0) {
        fence_after_sync();
        umma_tile(tm_W, uH1, 128, 1, uE, 128, 1, 0, 128, 8, 0);
        mma_commit(&bar_dwe);
      }
      // coalesced output: warp per row
#pragma unroll 1
      for (int rb = 0; rb < 32; rb += 8) {
        float4 old[8];
        int dst[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          dst[i] = __shfl_sync(0xffffffffu, my_dst, rb + i);
          old[i] = (dst[i] >= 0 && p.g_eout) ? ld_stream_f4(p.g_eout + (long long)dst[i] * H + 4 * lane)
                                             : make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float4 v = *reinterpret_cast<const float4*>(stage + stage_off(w * 32 + rb + i, 4 * lane));
          if (dst[i] >= 0)
            st_stream_f4(p.g_e + (long long)dst[i] * H + 4 * lane,
                         make_float4(old[i].x + v.x, old[i].y + v.y, old[i].z + v.z, old[i].w + v.w));
        }
      }
      // the dWe product reads sH1 / sE: wait before the next tile overwrites them
      mbar_wait(&bar_dwe, it & 1);
      named_bar(1, 128);
    }
    // ---------------- flush the resident weight gradients, bias and scale gradients
    fence_after_sync();
    flush_acc_rows(tm_dW4 + lane_off, p.grad + GB_E_W4, 128, t);
    flush_acc_rows(tm_dW3 + lane_off, p.grad + GB_E_W3, 128, t);
    flush_acc_rows(tm_dW2 + lane_off, p.grad + GB_E_W2, 128, t);
    const int cp = t & 63;
    atomicAdd(p.grad + GB_E_B4 + 2 * cp, db4a);
    atomicAdd(p.grad + GB_E_B4 + 2 * cp + 1, db4b);
    atomicAdd(p.grad + GB_E_B3 + 2 * cp, db3a);
    atomicAdd(p.grad + GB_E_B3 + 2 * cp + 1, db3b);
    atomicAdd(p.grad + GB_E_B2 + 2 * cp, db2a);
    atomicAdd(p.grad + GB_E_B2 + 2 * cp + 1, db2b);
#pragma unroll
    for (int k = 0; k < 4; ++k) atomicAdd(p.grad + GB_E_SC + 32 * k + lane, dsc[k]);
  }

  fence_before_sync();
  __syncthreads();
  if ((tid >> 5) == 0) tmem_dealloc<512>(tmem_slot);
}

// ===================================================================================== node
struct NodeBwdParams {
  long long N;
  const float* g_xout;      // [N,128]
  const __half* h1;         // fp16 hidden saves of node_fwd
  const __half* h2;
  const __half* h3;
  float* g_x;               // [N,128] out: g_xout + dz1' W1'[:, :128]
  float* g_aggr;            // [N,128] out: dz1' W1'[:, 128:]
  __half* dz1;              // [N,128] out (fp16) for node_wgrad1
  const __half* W;          // packed node tiles: W1' (2 slots), W2', W3', W4'
  const float* b4;
  const float* scale;
  float* grad;
};

__global__ void __launch_bounds__(128, 1) node_bwd_kernel(const NodeBwdParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  check_smem_alignment(smem);
  uint8_t* sW[2] = {smem, smem + SLOT};
  uint8_t* sH1 = smem + 2 * SLOT;
  uint8_t* sH2 = smem + 3 * SLOT;
  uint8_t* sH3 = smem + 4 * SLOT;
  uint8_t* sG = smem + 5 * SLOT;
  uint8_t* stage = sH3;   // sH3 + sG
  SmemTail* tail = reinterpret_cast<SmemTail*>(smem + 6 * SLOT);
  uint64_t* bar_w = &tail->bar[0];
  uint64_t& bar_mma = tail->bar[2];
  uint32_t& tmem_slot = tail->tmem_slot;
  const float* sb4 = tail->vec[3];
  const float* ssc = tail->vec[4];
  const int t = threadIdx.x, w = t >> 5, lane = t & 31;
  copy_vec128(tail->vec[3], p.b4, t, blockDim.x);
  copy_vec128(tail->vec[4], p.scale, t, blockDim.x);
  if (t == 0) {
    mbar_init(&bar_w[0], 1);
    mbar_init(&bar_w[1], 1);
    mbar_init(&bar_mma, 1);
    mbar_init_fence();
  }
  if (w == 0) tmem_alloc<512>(&tmem_slot);
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t lane_off = (uint32_t)(w * 32) << 16;
  const uint32_t tm_W = tmem_slot, tm_dW4 = tmem_slot + 128, tm_dW3 = tmem_slot + 256, tm_dW2 = tmem_slot + 384;
  const uint32_t uW[2] = {smem_u32(sW[0]), smem_u32(sW[1])};
  const uint32_t uH1 = smem_u32(sH1), uH2 = smem_u32(sH2), uH3 = smem_u32(sH3), uG = smem_u32(sG);
  const long long num_tiles = (p.N + TILE_ROWS - 1) / TILE_ROWS;
  const uint32_t my_tiles = (uint32_t)((num_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x);
  const uint32_t total_loads = 5 * my_tiles;
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
  if (t == 0) {
    issue_load(0);
    issue_load(1);
  }
  float dsc[4] = {0.f, 0.f, 0.f, 0.f};
  float db1a = 0.f, db1b = 0.f, db2a = 0.f, db2b = 0.f, db3a = 0.f, db3b = 0.f, db4a = 0.f, db4b = 0.f;
  uint32_t it = 0;
  for (long long tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
    const long long row0 = tile * TILE_ROWS;
    const uint32_t q0 = 5 * it;
    auto rf = [&](int r) -> long long { long long g = row0 + r; return g < p.N ? g : -1; };
    load_rows_f16_to_tile(sH1, 128, 0, p.h1, w * 32, lane, rf);
    load_rows_f16_to_tile(sH2, 128, 0, p.h2, w * 32, lane, rf);
    load_rows_f16_to_tile(sH3, 128, 0, p.h3, w * 32, lane, rf);
    load_rows_f32_to_tile_b<16>(sG, 128, 0, p.g_xout, w * 32, 32, lane, rf);
    tile_ready();
    // ---- R4: z4 = h3 W4'^T, RMSNorm backward
    if (t == 0) {
      wait_load(q0);
      fence_after_sync();
      umma_tile(tm_W, uH3, 128, 0, uW[q0 & 1], 128, 0, 0, 128, 8, 0);
      mma_commit(&bar_mma);
    }
    mma_wait();
    epi_rmsnorm_bwd(tm_W + lane_off, sb4, ssc, sG, t, lane, dsc);
    tile_ready();
    // ---- D4
    if (t == 0) {
      fence_after_sync();
      umma_tile(tm_W, uG, 128, 0, uW[q0 & 1], 128, 1, 0, 128, 8, 0);
      umma_tile(tm_dW4, uG, 128, 1, uH3, 128, 1, 0, 128, 8, it > 0);
      mma_commit(&bar_mma);
    }
    tile_colsum_pair(sG, t, db4a, db4b);
    mma_wait();
    if (t == 0) issue_load(q0 + 2);
    epi_relu_mask_inplace(tm_W + lane_off, sH3, t);
    tile_ready();
    // ---- D3
    if (t == 0) {
      wait_load(q0 + 1);
      fence_after_sync();
      umma_tile(tm_W, uH3, 128, 0, uW[(q0 + 1) & 1], 128, 1, 0, 128, 8, 0);
      umma_tile(tm_dW3, uH3, 128, 1, uH2, 128, 1, 0, 128, 8, it > 0);
      mma_commit(&bar_mma);
    }
    tile_colsum_pair(sH3, t, db3a, db3b);
    mma_wait();
    if (t == 0) issue_load(q0 + 3);
    epi_relu_mask_inplace(tm_W + lane_off, sH2, t);
    tile_ready();
    // ---- D2
    if (t == 0) {
      wait_load(q0 + 2);
      fence_after_sync();
      umma_tile(tm_W, uH2, 128, 0, uW[(q0 + 2) & 1], 128, 1, 0, 128, 8, 0);
      umma_tile(tm_dW2, uH2, 128, 1, uH1, 128, 1, 0, 128, 8, it > 0);
      mma_commit(&bar_mma);
    }
    tile_colsum_pair(sH2, t, db2a, db2b);
    mma_wait();
    if (t == 0) issue_load(q0 + 4);
    epi_relu_mask_inplace(tm_W + lane_off, sH1, t);     // sH1 <- dz1'
    tile_ready();
    // ---- D1a: g_x = g_xout + dz1' W1'[:, :128]
    if (t == 0) {
      wait_load(q0 + 3);
      fence_after_sync();
      umma_tile(tm_W, uH1, 128, 0, uW[(q0 + 3) & 1], 128, 1, 0, 128, 8, 0);
      mma_commit(&bar_mma);
    }
    tile_colsum_pair(sH1, t, db1a, db1b);
    {  // dz1' tile -> global fp16 (half-warp per row)
      const int hw = lane >> 4, q = lane & 15;
#pragma unroll 4
      for (int i = 0; i < 16; ++i) {
        const int r = w * 32 + 2 * i + hw;
        const long long g = row0 + r;
        if (g < p.N)
          *reinterpret_cast<uint4*>(p.dz1 + g * H + 8 * q) =
              *reinterpret_cast<const uint4*>(sH1 + tile_off(128, r, 8 * q));
      }
    }
    mma_wait();
    if (t == 0) issue_load(q0 + 5);
    acc_to_stage(tm_W + lane_off, stage, t);
    fence_before_sync();
    __syncthreads();
    // ---- D1b issued now (working accumulator free); g_x written meanwhile
    if (t == 0) {
      wait_load(q0 + 4);
      fence_after_sync();
      umma_tile(tm_W, uH1, 128, 0, uW[(q0 + 4) & 1], 128, 1, 0, 128, 8, 0);
      mma_commit(&bar_mma);
    }
#pragma unroll 1
    for (int rb = 0; rb < 32; rb += 8) {
      float4 old[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const long long g = row0 + w * 32 + rb + i;
        old[i] = g < p.N ? ld_stream_f4(p.g_xout + g * H + 4 * lane) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const long long g = row0 + w * 32 + rb + i;
        const float4 v = *reinterpret_cast<const float4*>(stage + stage_off(w * 32 + rb + i, 4 * lane));
        if (g < p.N)
          st_stream_f4(p.g_x + g * H + 4 * lane,
                       make_float4(old[i].x + v.x, old[i].y + v.y, old[i].z + v.z, old[i].w + v.w));
      }
    }
    mma_wait();
    if (t == 0) issue_load(q0 + 6);
    __syncthreads();    // all reads of the g_x staging are done
    acc_to_stage(tm_W + lane_off, stage, t);
    fence_before_sync();
    __syncthreads();
#pragma unroll 1
    for (int rb = 0; rb < 32; rb += 8) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const long long g = row0 + w * 32 + rb + i;
        const float4 v = *reinterpret_cast<const float4*>(stage + stage_off(w * 32 + rb + i, 4 * lane));
        if (g < p.N) st_stream_f4(p.g_aggr + g * H + 4 * lane, v);
      }
    }
    __syncthreads();
  }
  fence_after_sync();
  flush_acc_rows(tm_dW4 + lane_off, p.grad + GB_N_W4, 128, t);
  flush_acc_rows(tm_dW3 + lane_off, p.grad + GB_N_W3, 128, t);
  flush_acc_rows(tm_dW2 + lane_off, p.grad + GB_N_W2, 128, t);
  const int cp = t & 63;
  atomicAdd(p.grad + GB_N_B4 + 2 * cp, db4a);
  atomicAdd(p.grad + GB_N_B4 + 2 * cp + 1, db4b);
  atomicAdd(p.grad + GB_N_B3 + 2 * cp, db3a);
  atomicAdd(p.grad + GB_N_B3 + 2 * cp + 1, db3b);
  atomicAdd(p.grad + GB_N_B2 + 2 * cp, db2a);
  atomicAdd(p.grad + GB_N_B2 + 2 * cp + 1, db2b);
  atomicAdd(p.grad + GB_N_B1 + 2 * cp, db1a);
  atomicAdd(p.grad + GB_N_B1 + 2 * cp + 1, db1b);
#pragma unroll
  for (int k = 0; k < 4; ++k) atomicAdd(p.grad + GB_N_SC + 32 * k + lane, dsc[k]);
  fence_before_sync();
  __syncthreads();
  if (w == 0) tmem_dealloc<512>(tmem_slot);
}

// ===================================================================================== P backward
// gP [N,256] fp32 = [dP_i | dP_j]:  g_x += gP [W_i ; W_j] ; dW_i = dP_i^T x ; dW_j = dP_j^T x ; db1 = colsum dP_i
struct PBwdParams {
  long long N;
  const float* gP;
  const float* x;
  float* g_x;               // read-modify-write
  const __half* Wcat;       // [256 x 128] panel tile (64 KB)
  float* grad;
};

__global__ void __launch_bounds__(128, 1) p_bwd_kernel(const PBwdParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  check_smem_alignment(smem);
  uint8_t* sW = smem;                 // 64 KB
  uint8_t* sGP = smem + 2 * SLOT;     // 128 x 256 fp16 = 64 KB ; later the fp32 staging of g_x
  uint8_t* sX = smem + 4 * SLOT;      // 32 KB
  SmemTail* tail = reinterpret_cast<SmemTail*>(smem + 5 * SLOT);
  uint64_t& wbar = tail->bar[0];
  uint64_t& bar_mma = tail->bar[1];
  uint32_t& tmem_slot = tail->tmem_slot;
  const int t = threadIdx.x, w = t >> 5, lane = t & 31;
  if (t == 0) {
    mbar_init(&wbar, 1);
    mbar_init(&bar_mma, 1);
    mbar_init_fence();
  }
  if (w == 0) tmem_alloc<512>(&tmem_slot);
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t lane_off = (uint32_t)(w * 32) << 16;
  const uint32_t tm_W = tmem_slot, tm_dWi = tmem_slot + 128, tm_dWj = tmem_slot + 256;
  if (t == 0) {
    mbar_arrive_expect_tx(&wbar, 2 * SLOT);
    bulk_g2s(sW, p.Wcat, SLOT, &wbar);
    bulk_g2s(sW + SLOT, reinterpret_cast<const uint8_t*>(p.Wcat) + SLOT, SLOT, &wbar);
  }
  mbar_wait(&wbar, 0);
  const uint32_t uW = smem_u32(sW), uGP = smem_u32(sGP), uX = smem_u32(sX);
  const long long num_tiles = (p.N + TILE_ROWS - 1) / TILE_ROWS;
  uint32_t phase = 0, it = 0;
  float db1a = 0.f, db1b = 0.f;
  for (long long tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
    const long long row0 = tile * TILE_ROWS;
    auto rf = [&](int r) -> long long { long long g = row0 + r; return g < p.N ? g : -1; };
    // gP rows are 256 floats: two 128-column halves of a row live 128 floats apart
    load_rows_f32_to_tile_b<16>(sGP, 128, 0, p.gP, w * 32, 32, lane,
                                [&](int r) -> long long { long long g = row0 + r; return g < p.N ? 2 * g : -1; });
    load_rows_f32_to_tile_b<16>(sGP, 128, 128, p.gP, w * 32, 32, lane,
                                [&](int r) -> long long { long long g = row0 + r; return g < p.N ? 2 * g + 1 : -1; });
    load_rows_f32_to_tile_b<16>(sX, 128, 0, p.x, w * 32, 32, lane, rf);
    fence_async_smem();
    fence_before_sync();
    __syncthreads();
    if (t == 0) {
      fence_after_sync();
      // g_x part: [128 x 256] x [256 x 128]; B = MN-major view of the 256-row weight tile
      umma_tile(tm_W, uGP, 128, 0, uW, 256, 1, 0, 128, 16, 0);
      // dW_i, dW_j: A = MN-major views of the two column halves of the gP tile
      {
        const uint32_t idesc = make_idesc(128, 128, 1, 1);
        for (uint32_t ks = 0; ks < 8; ++ks) {
          mma_f16(tm_dWi, desc_mnmajor(uGP, 128, ks, 0), desc_mnmajor(uX, 128, ks, 0), idesc, (ks > 0 || it > 0) ? 1u : 0u);
        }
        for (uint32_t ks = 0; ks < 8; ++ks) {
          mma_f16(tm_dWj, desc_mnmajor(uGP, 128, ks, 128), desc_mnmajor(uX, 128, ks, 0), idesc, (ks > 0 || it > 0) ? 1u : 0u);
        }
      }
      mma_commit(&bar_mma);
    }
    tile_colsum_pair(sGP, t, db1a, db1b);       // columns 0..127 = dP_i
    mbar_wait(&bar_mma, phase);
    phase ^= 1;
    fence_after_sync();
    acc_to_stage(tm_W + lane_off, sGP, t);
    fence_before_sync();
    __syncthreads();
#pragma unroll 1
    for (int rb = 0; rb < 32; rb += 8) {
      float4 old[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const long long g = row0 + w * 32 + rb + i;
        old[i] = g < p.N ? ld_f4(p.g_x + g * H + 4 * lane) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const long long g = row0 + w * 32 + rb + i;
        const float4 v = *reinterpret_cast<const float4*>(sGP + stage_off(w * 32 + rb + i, 4 * lane));
        if (g < p.N)
          *reinterpret_cast<float4*>(p.g_x + g * H + 4 * lane) =
              make_float4(old[i].x + v.x, old[i].y + v.y, old[i].z + v.z, old[i].w + v.w);
      }
    }
    __syncthreads();
  }
  fence_after_sync();
  // dW_i -> edge_block.0.weight[:, 128:256], dW_j -> [:, 256:384]
  flush_acc_rows(tm_dWi + lane_off, p.grad + GB_E_W1 + 128, 384, t);
  flush_acc_rows(tm_dWj + lane_off, p.grad + GB_E_W1 + 256, 384, t);
  const int cp = t & 63;
  atomicAdd(p.grad + GB_E_B1 + 2 * cp, db1a);
  atomicAdd(p.grad + GB_E_B1 + 2 * cp + 1, db1b);
  fence_before_sync();
  __syncthreads();
  if (w == 0) tmem_dealloc<512>(tmem_slot);
}

// ===================================================================================== node layer-1 wgrad
// dW1' [128, 256] += dz1'^T [x | aggr]
struct NodeWgrad1Params {
  long long N;
  const __half* dz1;        // [N,128] fp16
  const float* x;
  const float* aggr;
  float* grad;
};

__global__ void __launch_bounds__(128, 1) node_wgrad1_kernel(const NodeWgrad1Params p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  check_smem_alignment(smem);
  uint8_t* sG = smem;                 // 32 KB
  uint8_t* sU = smem + SLOT;          // 128 x 256 fp16 = 64 KB
  SmemTail* tail = reinterpret_cast<SmemTail*>(smem + 3 * SLOT);
  uint64_t& bar_mma = tail->bar[0];
  uint32_t& tmem_slot = tail->tmem_slot;
  const int t = threadIdx.x, w = t >> 5, lane = t & 31;
  if (t == 0) {
    mbar_init(&bar_mma, 1);
    mbar_init_fence();
  }
  if (w == 0) tmem_alloc<256>(&tmem_slot);
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t lane_off = (uint32_t)(w * 32) << 16;
  const uint32_t tm = tmem_slot;
  const uint32_t uG = smem_u32(sG), uU = smem_u32(sU);
  const long long num_tiles = (p.N + TILE_ROWS - 1) / TILE_ROWS;
  uint32_t phase = 0, it = 0;
  for (long long tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
    const long long row0 = tile * TILE_ROWS;
    auto rf = [&](int r) -> long long { long long g = row0 + r; return g < p.N ? g : -1; };
    load_rows_f16_to_tile(sG, 128, 0, p.dz1, w * 32, lane, rf);
    load_rows_f32_to_tile_b<16>(sU, 128, 0, p.x, w * 32, 32, lane, rf);
    load_rows_f32_to_tile_b<16>(sU, 128, 128, p.aggr, w * 32, 32, lane, rf);
    fence_async_smem();
    fence_before_sync();
    __syncthreads();
    if (t == 0) {
      fence_after_sync();
      const uint32_t idesc = make_idesc(128, 128, 1, 1);
      for (uint32_t ks = 0; ks < 8; ++ks)
        mma_f16(tm, desc_mnmajor(uG, 128, ks, 0), desc_mnmajor(uU, 128, ks, 0), idesc, (ks > 0 || it > 0) ? 1u : 0u);
      for (uint32_t ks = 0; ks < 8; ++ks)
        mma_f16(tm + 128, desc_mnmajor(uG, 128, ks, 0), desc_mnmajor(uU, 128, ks, 128), idesc,
                (ks > 0 || it > 0) ? 1u : 0u);
      mma_commit(&bar_mma);
    }
    mbar_wait(&bar_mma, phase);
    phase ^= 1;
    fence_after_sync();
    __syncthreads();
  }
  flush_acc_rows(tm + lane_off, p.grad + GB_N_W1, 256, t);
  flush_acc_rows(tm + 128 + lane_off, p.grad + GB_N_W1 + 128, 256, t);
  fence_before_sync();
  __syncthreads();
  if (w == 0) tmem_dealloc<256>(tmem_slot);
}

// ===================================================================================== C ABI
static constexpr int EDGE_BWD_SMEM = 7 * SLOT + TAIL_BYTES;
static constexpr int NODE_BWD_SMEM = 6 * SLOT + TAIL_BYTES;
static constexpr int P_BWD_SMEM = 5 * SLOT + TAIL_BYTES;
static constexpr int NODE_WGRAD1_SMEM = 3 * SLOT + TAIL_BYTES;

extern "C" int64_t mgn_grad_block_floats(void) { return GB_FLOATS; }

// Backward of the edge update + aggregation (autograd of layers.py:687-698, 703-719).
// gP [N,256] is zeroed here and receives the gradients of the per-node partial products.
extern "C" int mgn_edge_bwd(int64_t E, int64_t N, const float* e_in, const int32_t* rowmap, const float* g_eout,
                            const float* g_aggr, const void* P, const int32_t* senders, const int32_t* receivers,
                            float* g_e, float* gP, const void* packed, const float* b2, const float* b3,
                            const float* b4, const float* scale, float* grad_block, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  MGN_CHECK_ARG(E >= 0 && N > 0 && gP, "edge_bwd: bad sizes");
  MGN_CUDA(cudaMemsetAsync(gP, 0, sizeof(float) * N * 256, stream));
  if (E == 0) return MGN_OK;
  MGN_CHECK_ARG(e_in && g_aggr && P && senders && receivers && g_e && packed && b2 && b3 && b4 && scale && grad_block,
                "edge_bwd: null pointer");
  EdgeBwdParams p{};
  p.E = E; p.e_in = e_in; p.rowmap = rowmap; p.g_eout = g_eout; p.g_aggr = g_aggr; p.P = (const __half*)P;
  p.senders = senders; p.receivers = receivers; p.g_e = g_e; p.gP = gP;
  p.W = reinterpret_cast<const __half*>(reinterpret_cast<const uint8_t*>(packed) + PK_EDGE);
  p.b2 = b2; p.b3 = b3; p.b4 = b4; p.scale = scale; p.grad = grad_block;
  static bool attr = false;
  if (!attr) {
    MGN_CUDA(cudaFuncSetAttribute(edge_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, EDGE_BWD_SMEM));
    attr = true;
  }
  const int grid = (int)std::min<int64_t>(mgn_num_sms(), ceil_div64(E, TILE_ROWS));
  edge_bwd_kernel<<<grid, 256, EDGE_BWD_SMEM, stream>>>(p);
  MGN_LAUNCH_CHECK();
  return MGN_OK;
}

// Backward of the node update (autograd of layers.py:733-746, 699) from the fp16 hidden saves.
extern "C" int mgn_node_bwd(int64_t N, const float* g_xout, const void* h1, const void* h2, const void* h3,
                            float* g_x, float* g_aggr, void* dz1, const void* packed, const float* b4,
                            const float* scale, float* grad_block, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  MGN_CHECK_ARG(N > 0 && g_xout && h1 && h2 && h3 && g_x && g_aggr && dz1 && packed && b4 && scale && grad_block,
                "node_bwd: bad args");
  NodeBwdParams p{};
  p.N = N; p.g_xout = g_xout; p.h1 = (const __half*)h1; p.h2 = (const __half*)h2; p.h3 = (const __half*)h3;
  p.g_x = g_x; p.g_aggr = g_aggr; p.dz1 = (__half*)dz1;
  p.W = reinterpret_cast<const __half*>(reinterpret_cast<const uint8_t*>(packed) + PK_NODE);
  p.b4 = b4; p.scale = scale; p.grad = grad_block;
  static bool attr = false;
  if (!attr) {
    MGN_CUDA(cudaFuncSetAttribute(node_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, NODE_BWD_SMEM));
    attr = true;
  }
  const int grid = (int)std::min<int64_t>(mgn_num_sms(), ceil_div64(N, TILE_ROWS));
  node_bwd_kernel<<<grid, 128, NODE_BWD_SMEM, stream>>>(p);
  MGN_LAUNCH_CHECK();
  return MGN_OK;
}

// Backward of the per-node partial products (mgn_p_fwd): g_x += gP [W_i ; W_j], dW_i, dW_j, db1.
extern "C" int mgn_p_bwd(int64_t N, const float* gP, const float* x, float* g_x, const void* packed,
                         float* grad_block, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  MGN_CHECK_ARG(N > 0 && gP && x && g_x && packed && grad_block, "p_bwd: bad args");
  PBwdParams p{N, gP, x, g_x,
               reinterpret_cast<const __half*>(reinterpret_cast<const uint8_t*>(packed) + PK_WCAT), grad_block};
  static bool attr = false;
  if (!attr) {
    MGN_CUDA(cudaFuncSetAttribute(p_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, P_BWD_SMEM));
    attr = true;
  }
  const int grid = (int)std::min<int64_t>(mgn_num_sms(), ceil_div64(N, TILE_ROWS));
  p_bwd_kernel<<<grid, 128, P_BWD_SMEM, stream>>>(p);
  MGN_LAUNCH_CHECK();
  return MGN_OK;
}

// Weight gradient of node_block layer 1: dW1' += dz1'^T [x | aggr].
extern "C" int mgn_node_wgrad1(int64_t N, const void* dz1, const float* x, const float* aggr, float* grad_block,
                               void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  MGN_CHECK_ARG(N > 0 && dz1 && x && aggr && grad_block, "node_wgrad1: bad args");
  NodeWgrad1Params p{N, (const __half*)dz1, x, aggr, grad_block};
  static bool attr = false;
  if (!attr) {
    MGN_CUDA(cudaFuncSetAttribute(node_wgrad1_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, NODE_WGRAD1_SMEM));
    attr = true;
  }
  const int grid = (int)std::min<int64_t>(mgn_num_sms(), ceil_div64(N, TILE_ROWS));
  node_wgrad1_kernel<<<grid, 128, NODE_WGRAD1_SMEM, stream>>>(p);
  MGN_LAUNCH_CHECK();
  return MGN_OK;
}
