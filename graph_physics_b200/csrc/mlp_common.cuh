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

}  // namespace mgn
