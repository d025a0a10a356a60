// mlp_v2.cuh -- thread mapping and per-thread epilogue pieces of the fused MLP kernels.
//
// Every kernel runs ONE 128-row tile at a time with 512 threads (16 warps).  Warp w owns TMEM
// lanes (= tile rows) 32*(w & 3) .. +31 -- the hardware's lane restriction -- and the 32-column
// quarter (w >> 2) of the 128-wide accumulator, so a thread's epilogue work is one
// tcgen05.ld.32x32b.x32: 32 columns of one row.  Sixteen warps per SM (four per scheduler) hide
// the shared-memory / TMEM / global latencies that four-warp tiles exposed (profiles/r01_*).
// Global traffic uses a second, coalesced mapping: warp w moves rows 8w .. 8w+7, a lane moves 16
// bytes of a 512-byte fp32 row (or half a warp moves a 256-byte fp16 row).
#pragma once
#include "mlp_common.cuh"

namespace mgn {

static constexpr int NT = 512;

struct Map {
  int tid, warp, lane;
  int lq, cq;          // lane quarter (rows 32*lq..), column quarter (columns 32*cq..)
  int row, c0;         // tile row / first column owned in the epilogues
  uint32_t lane_off;   // TMEM address offset of this warp's lanes
};
__device__ __forceinline__ Map make_map() {
  Map m;
  m.tid = threadIdx.x;
  m.warp = m.tid >> 5;
  m.lane = m.tid & 31;
  m.lq = m.warp & 3;
  m.cq = m.warp >> 2;
  m.row = m.lq * 32 + m.lane;
  m.c0 = m.cq * 32;
  m.lane_off = (uint32_t)(m.lq * 32) << 16;
  return m;
}

// dynamic shared memory tail (after the 32 KB slots): mbarriers, TMEM slot, row-reduction scratch
struct Tail2 {
  uint64_t bar[8];
  uint32_t tmem_slot;
  uint32_t pad[3];
  float red[4 * 128];
};
static_assert(sizeof(Tail2) <= TAIL_BYTES, "Tail2 too large");

// 32 consecutive floats of a global vector (bias / scale): every lane reads the same 128 bytes
__device__ __forceinline__ void load_vec32(const float* __restrict__ g, int c0, float4 (&b)[8]) {
#pragma unroll
  for (int j = 0; j < 8; ++j) b[j] = ldg_f4(g + c0 + 4 * j);
}

// sum of `partial` over the four column quarters of each row (all NT threads call)
__device__ __forceinline__ float row_sum4(float partial, float* red, const Map& m) {
  red[m.cq * 128 + m.row] = partial;
  __syncthreads();
  const float s = (red[m.row] + red[128 + m.row]) + (red[256 + m.row] + red[384 + m.row]);
  __syncthreads();
  return s;
}

// hidden layer: tile[row, c0..c0+32) <- fp16(relu(acc + bias)); optional fp16 copy to save[grow, ...]
__device__ __forceinline__ void epi_bias_relu_q(uint32_t taddr, const float4 (&bb)[8], uint8_t* tile, const Map& m,
                                                __half* save, long long grow) {
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
  store_row_chunk_f16(tile, 128, m.row, m.c0, v);
  if (save != nullptr) {
#pragma unroll
    for (int j = 0; j < 4; ++j)
      *reinterpret_cast<uint4*>(save + grow * H + m.c0 + 8 * j) =
          *reinterpret_cast<const uint4*>(tile + tile_off(128, m.row, m.c0 + 8 * j));
  }
}

// 32 fp32 of one row -> 16 packed fp16 pairs -> this thread's 16 columns of a TMEM-resident A operand (tc05.cuh
// mma_f16_ts: K element 2c in the low half of column c); optional fp16 copy of the row chunk to save[grow, c0..c0+32)
__device__ __forceinline__ void store_row_chunk_tmem(uint32_t taddr_a, const float (&v)[32], __half* save,
                                                     long long grow, int c0) {
  uint32_t r[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) r[j] = pack_h2(v[2 * j], v[2 * j + 1]);
  tmem_st16(taddr_a, r);
  if (save != nullptr) {
#pragma unroll
    for (int j = 0; j < 4; ++j)
      *reinterpret_cast<uint4*>(save + grow * H + c0 + 8 * j) = make_uint4(r[4 * j], r[4 * j + 1], r[4 * j + 2], r[4 * j + 3]);
  }
}
// hidden layer kept in tensor memory: A-operand columns <- fp16(relu(acc + bias))
__device__ __forceinline__ void epi_bias_relu_tmem_q(uint32_t taddr, uint32_t taddr_a, const float4 (&bb)[8],
                                                     const Map& m, __half* save, long long grow) {
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
  store_row_chunk_tmem(taddr_a, v, save, grow, m.c0);
}

// ReLU backward in place: tile holds h = relu(z); overwrite with fp16(acc * [h > 0]).  The mask is applied to the
// PACKED result (h > 0 per half -> 0xFFFF / 0 by one HSET2 per pair, one AND): no fp16 -> fp32 unpacking of h, no
// per-element select; same bits as fp16(h > 0 ? acc : 0).
__device__ __forceinline__ void epi_relu_mask_q(uint32_t taddr, uint8_t* tile, const Map& m) {
  float v[32];
  tmem_ld32(taddr, v);
  uint4 hw[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) hw[j] = *reinterpret_cast<const uint4*>(tile + tile_off(128, m.row, m.c0 + 8 * j));
  tmem_ld_wait();
  const __half2 zero2 = __float2half2_rn(0.f);
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const uint32_t h[4] = {hw[j].x, hw[j].y, hw[j].z, hw[j].w};
    uint32_t o[4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
      o[i] = pack_h2(v[8 * j + 2 * i], v[8 * j + 2 * i + 1]) & __hgt2_mask(*reinterpret_cast<const __half2*>(&h[i]), zero2);
    *reinterpret_cast<uint4*>(tile + tile_off(128, m.row, m.c0 + 8 * j)) = make_uint4(o[0], o[1], o[2], o[3]);
  }
}

// accumulator quarter -> fp32 staging tile (stage_off layout, 64 KB)
__device__ __forceinline__ void acc_to_stage_q(uint32_t taddr, uint8_t* stage, const Map& m) {
  float v[32];
  tmem_ld32(taddr, v);
  tmem_ld_wait();
#pragma unroll
  for (int j = 0; j < 8; ++j)
    *reinterpret_cast<float4*>(stage + stage_off(m.row, m.c0 + 4 * j)) =
        make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
}

// accumulator quarter -> global with red.add (row pitch `ld` floats)
__device__ __forceinline__ void flush_acc_q(uint32_t taddr, float* __restrict__ dst, int ld, const Map& m) {
  float v[32];
  tmem_ld32(taddr, v);
  tmem_ld_wait();
#pragma unroll
  for (int j = 0; j < 8; ++j)
    red_add_f4(dst + (long long)m.row * ld + m.c0 + 4 * j, make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]));
}

// column sums of a 128 x 128 fp16 tile: thread adds rows [16*(tid>>6), +16) of columns 2*(tid&63), +1
__device__ __forceinline__ void tile_colsum_q(const uint8_t* tile, int tid, float& a0, float& a1) {
  const int cp = tid & 63, rbase = (tid >> 6) * 16;
#pragma unroll
  for (int r = 0; r < 16; ++r) {
    const uint32_t u = *reinterpret_cast<const uint32_t*>(tile + tile_off(128, rbase + r, 2 * cp));
    const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&u));
    a0 += f.x;
    a1 += f.y;
  }
}

// v[32] = 32 consecutive fp32 columns (c0..) of tile row m.row -> LINEAR fp32 tile [128][128] (512-byte rows, the
// layout bulk copies / bulk reductions to global need).  A plain row-owner store would be a 32-way bank conflict
// (row pitch 512 B), so lane l writes its 16-byte chunk c to slot c ^ (l & 7); the registers are pre-permuted by a
// three-stage conditional butterfly (no dynamic register indexing).  v is clobbered.
__device__ __forceinline__ void store_row32_linear(uint8_t* tile, const Map& m, float (&v)[32]) {
  const int k = m.lane & 7;
#pragma unroll
  for (int b = 1; b < 8; b <<= 1) {
    const bool sw = (k & b) != 0;
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      if (c & b) continue;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float lo = v[4 * c + i], hi = v[4 * (c ^ b) + i];
        v[4 * c + i] = sw ? hi : lo;
        v[4 * (c ^ b) + i] = sw ? lo : hi;
      }
    }
  }
  uint8_t* base = tile + m.row * 512 + m.c0 * 4;
#pragma unroll
  for (int c = 0; c < 8; ++c)
    *reinterpret_cast<float4*>(base + 16 * (c ^ k)) = make_float4(v[4 * c], v[4 * c + 1], v[4 * c + 2], v[4 * c + 3]);
}
// same, adding to what the linear tile already holds (the upstream rows a bulk copy put there)
__device__ __forceinline__ void add_row32_linear(uint8_t* tile, const Map& m, float (&v)[32]) {
  const int k = m.lane & 7;
#pragma unroll
  for (int b = 1; b < 8; b <<= 1) {
    const bool sw = (k & b) != 0;
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      if (c & b) continue;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float lo = v[4 * c + i], hi = v[4 * (c ^ b) + i];
        v[4 * c + i] = sw ? hi : lo;
        v[4 * (c ^ b) + i] = sw ? lo : hi;
      }
    }
  }
  uint8_t* base = tile + m.row * 512 + m.c0 * 4;
#pragma unroll
  for (int c = 0; c < 8; ++c) {
    float4* ptr = reinterpret_cast<float4*>(base + 16 * (c ^ k));
    const float4 o = *ptr;
    *ptr = make_float4(o.x + v[4 * c], o.y + v[4 * c + 1], o.z + v[4 * c + 2], o.w + v[4 * c + 3]);
  }
}

// Resident 128 x 128 fp32 accumulator (all 16 warps' quarters) += into a [128][ld] fp32 global matrix: rows pass through a
// row-linear 64 KB staging tile and leave as bulk reductions performed at L2 (one 64 KB copy when the destination rows
// are contiguous, else one 512-byte copy per row) -- a row-owner red.add touches 32 distinct lines per instruction and
// costs ~5 k cycles of the load/store path per matrix.  All NT threads call; `lin32` must be dead; ends with a barrier
// after the copies have read it.
__device__ __forceinline__ void flush_acc_bulk(uint32_t taddr, float* __restrict__ dst, int ld, uint8_t* lin32,
                                               const Map& m, uint64_t pol) {
  float v[32];
  tmem_ld32(taddr, v);
  tmem_ld_wait();
  store_row32_linear(lin32, m, v);
  fence_async_smem();
  __syncthreads();
  if (ld == 128) {
    if (m.tid == 0) {
      bulk_red_add_f32(dst, lin32, 128 * 512, pol);
      bulk_commit();
      bulk_wait_read0();
    }
  } else if (m.tid < 128) {
    bulk_red_add_f32(dst + (long long)m.tid * ld, lin32 + m.tid * 512, 512, pol);
    bulk_commit();
    bulk_wait_read0();
  }
  __syncthreads();
}

// Three resident accumulators at once (the end-of-kernel flush of dW4, dW3, dW2): every tile slot is dead by then, so
// the three are staged side by side (3 x 64 KB from `base`) and leave as three back-to-back bulk reductions -- one
// barrier pair and one read-out wait instead of three.  Contiguous destinations ([128][128] blocks).
__device__ __forceinline__ void flush_acc_bulk3(uint32_t t0, uint32_t t1, uint32_t t2, float* __restrict__ d0,
                                                float* __restrict__ d1, float* __restrict__ d2, uint8_t* base,
                                                const Map& m, uint64_t pol) {
  const uint32_t ta[3] = {t0, t1, t2};
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    float v[32];
    tmem_ld32(ta[i], v);
    tmem_ld_wait();
    store_row32_linear(base + i * 65536, m, v);
  }
  fence_async_smem();
  __syncthreads();
  if (m.tid == 0) {
    bulk_red_add_f32(d0, base, 65536, pol);
    bulk_red_add_f32(d1, base + 65536, 65536, pol);
    bulk_red_add_f32(d2, base + 2 * 65536, 65536, pol);
    bulk_commit();
    bulk_wait_read0();
  }
  __syncthreads();
}

// ReLU backward in place (tile: h -> fp16(acc * [h > 0])) plus an fp32 copy of the result in a linear tile
__device__ __forceinline__ void epi_relu_mask_q32(uint32_t taddr, uint8_t* tile, uint8_t* lin32, const Map& m) {
  float v[32], h[32];
  tmem_ld32(taddr, v);
  load_row_chunk_f16(tile, 128, m.row, m.c0, h);
  tmem_ld_wait();
#pragma unroll
  for (int j = 0; j < 32; ++j) v[j] = h[j] > 0.f ? v[j] : 0.f;
  store_row_chunk_f16(tile, 128, m.row, m.c0, v);
  store_row32_linear(lin32, m, v);
}

// ---- coalesced movers: warp w handles tile rows 8w .. 8w+7; `grow(i)` = global row of the i-th or -1
template <class RowFn>
__device__ __forceinline__ void load8_f32_to_tile(uint8_t* tile, uint32_t tile_rows, uint32_t col0,
                                                  const float* __restrict__ src, const Map& m, RowFn grow,
                                                  uint64_t pol) {
  float4 v[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const long long g = grow(i);
    v[i] = (g >= 0) ? ld_f4_hint(src + g * H + 4 * m.lane, pol) : make_float4(0.f, 0.f, 0.f, 0.f);
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    uint2 pk;
    pk.x = pack_h2(v[i].x, v[i].y);
    pk.y = pack_h2(v[i].z, v[i].w);
    *reinterpret_cast<uint2*>(tile + tile_off(tile_rows, m.warp * 8 + i, col0 + 4 * m.lane)) = pk;
  }
}
// the same through L2 (ld.global.cg), for rows a chained producer launch may still have been writing when this launch
// started (tile-level chaining, mlp_common.cuh)
template <class RowFn>
__device__ __forceinline__ void load8_f32_to_tile_cg(uint8_t* tile, uint32_t tile_rows, uint32_t col0,
                                                     const float* __restrict__ src, const Map& m, RowFn grow) {
  float4 v[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const long long g = grow(i);
    v[i] = (g >= 0) ? __ldcg(reinterpret_cast<const float4*>(src + g * H + 4 * m.lane)) : make_float4(0.f, 0.f, 0.f, 0.f);
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    uint2 pk;
    pk.x = pack_h2(v[i].x, v[i].y);
    pk.y = pack_h2(v[i].z, v[i].w);
    *reinterpret_cast<uint2*>(tile + tile_off(tile_rows, m.warp * 8 + i, col0 + 4 * m.lane)) = pk;
  }
}
// tile row <- fp16(a[ga(i)] + b[gb(i)]).  Consecutive rows that name the same b row (edges of one receiver, CSR
// order) reuse the registers of the previous row instead of fetching it again (warp-uniform test).
template <class RowFnA, class RowFnB>
__device__ __forceinline__ void load8_sum2_to_tile(uint8_t* tile, const float* __restrict__ a,
                                                   const float* __restrict__ b, const Map& m, RowFnA ga, RowFnB gb,
                                                   uint64_t pol_a, uint64_t pol_b) {
  float4 va[8], vb[8];
  long long prev = -2;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const long long ia = ga(i), ib = gb(i);
    va[i] = (ia >= 0) ? ld_f4_hint(a + ia * H + 4 * m.lane, pol_a) : make_float4(0.f, 0.f, 0.f, 0.f);
    if (i > 0 && ib == prev) vb[i] = vb[i - 1];
    else vb[i] = (ib >= 0) ? ld_f4_hint(b + ib * H + 4 * m.lane, pol_b) : make_float4(0.f, 0.f, 0.f, 0.f);
    prev = ib;
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    uint2 pk;
    pk.x = pack_h2(va[i].x + vb[i].x, va[i].y + vb[i].y);
    pk.y = pack_h2(va[i].z + vb[i].z, va[i].w + vb[i].w);
    *reinterpret_cast<uint2*>(tile + tile_off(128, m.warp * 8 + i, 4 * m.lane)) = pk;
  }
}
// fp16 [*,128] global rows -> tile: half-warp per row, 4 x 2 rows per warp
template <class RowFn>
__device__ __forceinline__ void load8_f16_to_tile(uint8_t* tile, const __half* __restrict__ src, const Map& m,
                                                  RowFn grow) {
  const int hw = m.lane >> 4, q = m.lane & 15;
  uint4 v[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const long long g = grow(2 * i + hw);
    v[i] = (g >= 0) ? ldg_u4(src + g * H + 8 * q) : make_uint4(0, 0, 0, 0);
  }
#pragma unroll
  for (int i = 0; i < 4; ++i)
    *reinterpret_cast<uint4*>(tile + tile_off(128, m.warp * 8 + 2 * i + hw, 8 * q)) = v[i];
}

// the same through L2 (ld.global.cg): for rows another kernel may have written while this one was already running
// (tile-level chaining, mlp_common.cuh) -- the non-coherent path of __ldg is for data that is read-only for the launch
template <class RowFn>
__device__ __forceinline__ void load8_f16_to_tile_cg(uint8_t* tile, const __half* __restrict__ src, const Map& m,
                                                     RowFn grow) {
  const int hw = m.lane >> 4, q = m.lane & 15;
  uint4 v[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const long long g = grow(2 * i + hw);
    v[i] = (g >= 0) ? __ldcg(reinterpret_cast<const uint4*>(src + g * H + 8 * q)) : make_uint4(0, 0, 0, 0);
  }
#pragma unroll
  for (int i = 0; i < 4; ++i)
    *reinterpret_cast<uint4*>(tile + tile_off(128, m.warp * 8 + 2 * i + hw, 8 * q)) = v[i];
}

}  // namespace mgn
