// graph_world.cu -- world edges: all OBSTACLE <-> NORMAL node pairs within `radius` in world space.
// Replaces the cKDTree.query_pairs + node-type masks of add_world_edges
// (graphphysics/dataset/preprocessing.py:92-140) with a uniform cell list on the device:
//   mgn_world_cell_keys  : key[n] = ((sample * DX + cx) * DY + cy) * DZ + cz, cell edge = 1.001 * radius
//   (caller sorts the keys -- plumbing -- and passes keys_sorted / order back)
//   mgn_world_pairs      : one warp per first-set node; the 3 x 3 neighbouring (cx, cy) columns are 9 contiguous
//                          key ranges [cz-1 .. cz+1]; count pass, then fill pass at the caller's exclusive offsets.
// The distance test is the one scipy makes: squared Euclidean distance of the float64-converted coordinates <= r^2.
// Pair order is irrelevant: the caller runs to_undirected (sort + unique) on pairs ++ mesh edges, as the reference does.
#include <algorithm>

#include "common.cuh"

static constexpr int TPB = 256;

struct WorldGrid {
  double origin[3];
  double inv_cell;
  long long dims[3];
};

__device__ __forceinline__ void cell_of(const float* __restrict__ pos, int dim, int64_t n, const WorldGrid& g,
                                        long long c[3]) {
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    double v = a < dim ? ((double)pos[n * dim + a] - g.origin[a]) * g.inv_cell : 0.0;
    long long q = (long long)floor(v);
    c[a] = q < 0 ? 0 : (q >= g.dims[a] ? g.dims[a] - 1 : q);
  }
}

__global__ void world_keys_kernel(const float* __restrict__ pos, int dim, int64_t N, const int32_t* __restrict__ sample,
                                  WorldGrid g, int64_t* __restrict__ keys) {
  for (int64_t n = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; n < N; n += (int64_t)gridDim.x * blockDim.x) {
    long long c[3];
    cell_of(pos, dim, n, g, c);
    long long b = sample ? sample[n] : 0;
    keys[n] = ((b * g.dims[0] + c[0]) * g.dims[1] + c[1]) * g.dims[2] + c[2];
  }
}

__device__ __forceinline__ int64_t lower_bound64(const int64_t* __restrict__ a, int64_t n, int64_t v) {
  int64_t lo = 0, hi = n;
  while (lo < hi) {
    int64_t mid = (lo + hi) >> 1;
    if (a[mid] < v) lo = mid + 1; else hi = mid;
  }
  return lo;
}

template <bool FILL>
__global__ void world_pairs_kernel(const float* __restrict__ pos, int dim, int64_t N, const uint8_t* __restrict__ cls,
                                   const int32_t* __restrict__ sample, const int64_t* __restrict__ keys_sorted,
                                   const int32_t* __restrict__ order, WorldGrid g, double r2,
                                   int64_t* __restrict__ counts, const int64_t* __restrict__ offs,
                                   int64_t* __restrict__ pairs, int64_t M) {
  const int lane = threadIdx.x & 31;
  const int64_t warp0 = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t i = warp0; i < N; i += nwarps) {
    if (cls[i] != 1) {
      if (!FILL && lane == 0) counts[i] = 0;
      continue;
    }
    long long c[3];
    cell_of(pos, dim, i, g, c);
    const long long b = sample ? sample[i] : 0;
    double pi[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) pi[a] = a < dim ? (double)pos[i * dim + a] : 0.0;
    int64_t found = 0;
    int64_t out = FILL ? offs[i] : 0;
    for (int dx = -1; dx <= 1; ++dx) {
      const long long cx = c[0] + dx;
      if (cx < 0 || cx >= g.dims[0]) continue;
      for (int dy = -1; dy <= 1; ++dy) {
        const long long cy = c[1] + dy;
        if (cy < 0 || cy >= g.dims[1]) continue;
        const long long z0 = c[2] > 0 ? c[2] - 1 : 0;
        const long long z1 = c[2] + 1 < g.dims[2] ? c[2] + 1 : g.dims[2] - 1;
        const long long base = ((b * g.dims[0] + cx) * g.dims[1] + cy) * g.dims[2];
        const int64_t lo = lower_bound64(keys_sorted, N, base + z0);
        const int64_t hi = lower_bound64(keys_sorted, N, base + z1 + 1);
        for (int64_t q0 = lo; q0 < hi; q0 += 32) {
          const int64_t q = q0 + lane;
          bool hit = false;
          int64_t j = -1;
          if (q < hi) {
            j = order[q];
            if (cls[j] == 2) {
              double d2 = 0.0;
#pragma unroll
              for (int a = 0; a < 3; ++a) {
                if (a < dim) {
                  const double d = pi[a] - (double)pos[j * dim + a];
                  d2 += d * d;
                }
              }
              hit = d2 <= r2;
            }
          }
          const unsigned m = __ballot_sync(0xffffffffu, hit);
          if (FILL && hit) {
            const int64_t o = out + __popc(m & ((1u << lane) - 1u));
            pairs[o] = i;
            pairs[M + o] = j;
          }
          out += __popc(m);
          found += __popc(m);
        }
      }
    }
    if (!FILL && lane == 0) counts[i] = found;
  }
}

static int make_grid(const double* origin, double radius, const int64_t* dims, WorldGrid* g) {
  MGN_CHECK_ARG(radius > 0.0, "mgn_world: radius must be positive");
  for (int a = 0; a < 3; ++a) {
    MGN_CHECK_ARG(dims[a] >= 1, "mgn_world: grid dims must be >= 1");
    g->origin[a] = origin[a];
    g->dims[a] = dims[a];
  }
  g->inv_cell = 1.0 / (1.001 * radius);
  return MGN_OK;
}

// dims[a] the caller must pass: floor((max_a - origin_a) / (1.001 radius)) + 1 (>= 1); origin = per-axis minimum.
extern "C" int mgn_world_cell_keys(const float* pos, int dim, int64_t N, const int32_t* sample, const double* origin,
                                   double radius, const int64_t* dims, int64_t* keys, void* stream_) {
  MGN_CHECK_ARG(dim == 2 || dim == 3, "mgn_world_cell_keys: dim must be 2 or 3");
  WorldGrid g;
  int rc = make_grid(origin, radius, dims, &g);
  if (rc) return rc;
  if (N == 0) return MGN_OK;
  cudaStream_t st = (cudaStream_t)stream_;
  int blocks = (int)std::min<int64_t>(ceil_div64(N, TPB), mgn_num_sms() * 8);
  world_keys_kernel<<<blocks, TPB, 0, st>>>(pos, dim, N, sample, g, keys);
  MGN_LAUNCH_CHECK();
  return MGN_OK;
}

// offs == NULL: count pass (counts[N] written).  offs != NULL: fill pass, pairs [2,M] int64 (row 0 = first-set node).
extern "C" int mgn_world_pairs(const float* pos, int dim, int64_t N, const uint8_t* cls, const int32_t* sample,
                               const int64_t* keys_sorted, const int32_t* order, const double* origin, double radius,
                               const int64_t* dims, int64_t* counts, const int64_t* offs, int64_t* pairs, int64_t M,
                               void* stream_) {
  MGN_CHECK_ARG(dim == 2 || dim == 3, "mgn_world_pairs: dim must be 2 or 3");
  WorldGrid g;
  int rc = make_grid(origin, radius, dims, &g);
  if (rc) return rc;
  if (N == 0) return MGN_OK;
  cudaStream_t st = (cudaStream_t)stream_;
  int blocks = (int)std::min<int64_t>(ceil_div64(N * 32, TPB), mgn_num_sms() * 8);
  const double r2 = radius * radius;
  if (offs == nullptr) {
    MGN_CHECK_ARG(counts != nullptr, "mgn_world_pairs: counts is NULL");
    world_pairs_kernel<false><<<blocks, TPB, 0, st>>>(pos, dim, N, cls, sample, keys_sorted, order, g, r2, counts,
                                                      nullptr, nullptr, 0);
  } else {
    MGN_CHECK_ARG(pairs != nullptr || M == 0, "mgn_world_pairs: pairs is NULL");
    if (M == 0) return MGN_OK;
    world_pairs_kernel<true><<<blocks, TPB, 0, st>>>(pos, dim, N, cls, sample, keys_sorted, order, g, r2, nullptr,
                                                     offs, pairs, M);
  }
  MGN_LAUNCH_CHECK();
  return MGN_OK;
}
