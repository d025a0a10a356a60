// packed weight layout of one GraphNetBlock (fp16 panel tiles, see tc05.cuh)
#pragma once
#include <stdint.h>
#include <cuda_runtime.h>
// second half of the deterministic receiver aggregation (edge_fwd_hl.cu)
int mgn_launch_aggr_fixup(int64_t E, int64_t N, int slot_rows, const int32_t* receivers, const int32_t* rowptr,
                          const float* carry, float* aggr, cudaStream_t stream);
namespace mgn {
static constexpr int64_t PK_EDGE = 0;                 // W1e, W2, W3, W4   (4 x 32 KB)
static constexpr int64_t PK_WCAT = 4 * 32768;         // [W_i ; W_j] 256 x 128 (64 KB)
static constexpr int64_t PK_NODE = 6 * 32768;         // W1' 128 x 256 (64 KB), W2', W3', W4'
static constexpr int64_t PK_BYTES = 11 * 32768;       // 360,448 B
}  // namespace mgn

// flat fp32 gradient block of one GraphNetBlock (floats), in state-dict order:
// edge_block.{0,2,4,6}.{weight,bias}, edge_block.7.scale, node_block.* likewise
namespace mgn {
static constexpr int64_t GB_E_W1 = 0;                       // [128,384]
static constexpr int64_t GB_E_B1 = GB_E_W1 + 128 * 384;
static constexpr int64_t GB_E_W2 = GB_E_B1 + 128;
static constexpr int64_t GB_E_B2 = GB_E_W2 + 128 * 128;
static constexpr int64_t GB_E_W3 = GB_E_B2 + 128;
static constexpr int64_t GB_E_B3 = GB_E_W3 + 128 * 128;
static constexpr int64_t GB_E_W4 = GB_E_B3 + 128;
static constexpr int64_t GB_E_B4 = GB_E_W4 + 128 * 128;
static constexpr int64_t GB_E_SC = GB_E_B4 + 128;
static constexpr int64_t GB_N_W1 = GB_E_SC + 128;           // [128,256]
static constexpr int64_t GB_N_B1 = GB_N_W1 + 128 * 256;
static constexpr int64_t GB_N_W2 = GB_N_B1 + 128;
static constexpr int64_t GB_N_B2 = GB_N_W2 + 128 * 128;
static constexpr int64_t GB_N_W3 = GB_N_B2 + 128;
static constexpr int64_t GB_N_B3 = GB_N_W3 + 128 * 128;
static constexpr int64_t GB_N_W4 = GB_N_B3 + 128;
static constexpr int64_t GB_N_B4 = GB_N_W4 + 128 * 128;
static constexpr int64_t GB_N_SC = GB_N_B4 + 128;
static constexpr int64_t GB_FLOATS = GB_N_SC + 128;         // 181,504
}  // namespace mgn
