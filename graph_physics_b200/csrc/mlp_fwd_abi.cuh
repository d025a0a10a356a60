// packed weight layout of one GraphNetBlock (fp16 panel tiles, see tc05.cuh)
#pragma once
#include <stdint.h>
namespace mgn {
static constexpr int64_t PK_EDGE = 0;                 // W1e, W2, W3, W4   (4 x 32 KB)
static constexpr int64_t PK_WCAT = 4 * 32768;         // [W_i ; W_j] 256 x 128 (64 KB)
static constexpr int64_t PK_NODE = 6 * 32768;         // W1' 128 x 256 (64 KB), W2', W3', W4'
static constexpr int64_t PK_BYTES = 11 * 32768;       // 360,448 B
}  // namespace mgn
