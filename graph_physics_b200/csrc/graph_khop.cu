// graph_khop.cu -- K-hop expansion, MeshMask edge filtering and raw edge features.
// Replaces compute_k_hop_edge_index (graphphysics/utils/torch_graph.py:16-53),
// filter_edges (graphphysics/utils/meshmask.py:9-37) and add_edge_features
// (graphphysics/dataset/preprocessing.py:16-23 = T.Cartesian + T.Distance, norm=False).
#include <cub/cub.cuh>

#include "common.cuh"

static constexpr int TPB = 256;
static inline int blocks_for(int64_t n) { return (int)std::min<int64_t>(ceil_div64(n, TPB), 148 * 16); }

// row pointer of a coalesced (row-sorted) edge_index: rowptr[n] = lower_bound(row, n)
__global__ void row_ptr_kernel(const int64_t* __restrict__ row, int64_t E, int64_t N, int64_t* __restrict__ rowptr) {
  for (int64_t n = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; n <= N; n += (int64_t)gridDim.x * blockDim.x) {
    int64_t lo = 0, hi = E;
    while (lo < hi) {
      int64_t mid = (lo + hi) >> 1;
      if (row[mid] < n) lo = mid + 1; else hi = mid;
    }
    rowptr[n] = lo;
  }
}
// fan[k] = out-degree in A of col_k[k]  (number of 2-step candidates of entry k of A_k)
__global__ void khop_fan_kernel(const int64_t* __restrict__ eik, int64_t Ek, const int64_t* __restrict__ rowptr,
                                int64_t* __restrict__ fan) {
  for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < Ek; k += (int64_t)gridDim.x * blockDim.x) {
    int64_t j = eik[Ek + k];
    fan[k] = rowptr[j + 1] - rowptr[j];
  }
}
// candidates (i, l) for every (i, j) in A_k and (j, l) in A; self-loops (i == l) are
// replaced by the entry's own key (i, j), which the later unique pass drops.
__global__ void khop_expand_kernel(const int64_t* __restrict__ eik, int64_t Ek, const int64_t* __restrict__ ei,
                                   int64_t E, const int64_t* __restrict__ rowptr,
                                   const int64_t* __restrict__ offs, int64_t N, int64_t* __restrict__ keys) {
  for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < Ek; k += (int64_t)gridDim.x * blockDim.x) {
    int64_t i = eik[k], j = eik[Ek + k];
    int64_t own = i * N + j;
    int64_t o = offs[k];
    for (int64_t q = rowptr[j]; q < rowptr[j + 1]; ++q) {
      int64_t l = ei[E + q];
      keys[o++] = (l == i) ? own : i * N + l;
    }
    keys[offs[Ek] + k] = own;   // A_k itself (the "A_k +" term)
  }
}

extern "C" int64_t mgn_khop_workspace_bytes(int64_t E, int64_t Ek, int64_t num_nodes) {
  size_t a = 0;
  cub::DeviceScan::ExclusiveSum(nullptr, a, (const int64_t*)nullptr, (int64_t*)nullptr, Ek + 1);
  return (int64_t)((a + 255) / 256 * 256) + 8 * (num_nodes + 1 + 2 * (Ek + 1)) + 1024;
}

// Phase 1: number of candidate keys of one hop step (= sum of fan + Ek). Synchronises.
// `edge_index` [2,E] is the 1-hop adjacency A, `edge_index_k` [2,Ek] the current A_k;
// both must be coalesced (ascending (row, col)).  The workspace keeps rowptr/offsets
// for phase 2.
extern "C" int mgn_khop_count(const int64_t* edge_index, int64_t E, const int64_t* edge_index_k, int64_t Ek,
                              int64_t num_nodes, void* workspace, int64_t workspace_bytes, int64_t* total_host,
                              void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  MGN_CHECK_ARG(edge_index && edge_index_k && workspace && total_host && E > 0 && Ek > 0, "khop_count: bad args");
  MGN_CHECK_ARG(workspace_bytes >= mgn_khop_workspace_bytes(E, Ek, num_nodes), "khop_count: workspace too small");
  int64_t* rowptr = reinterpret_cast<int64_t*>(workspace);
  int64_t* fan = rowptr + (num_nodes + 1);
  int64_t* offs = fan + (Ek + 1);
  void* tmp = reinterpret_cast<void*>(((uintptr_t)(offs + (Ek + 1)) + 255) & ~(uintptr_t)255);
  size_t tmp_bytes = (size_t)workspace_bytes - ((uint8_t*)tmp - (uint8_t*)workspace);
  row_ptr_kernel<<<blocks_for(num_nodes + 1), TPB, 0, stream>>>(edge_index, E, num_nodes, rowptr);
  MGN_LAUNCH_CHECK();
  MGN_CUDA(cudaMemsetAsync(fan + Ek, 0, 8, stream));
  khop_fan_kernel<<<blocks_for(Ek), TPB, 0, stream>>>(edge_index_k, Ek, rowptr, fan);
  MGN_LAUNCH_CHECK();
  MGN_CUDA(cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, fan, offs, (int)(Ek + 1), stream));
  int64_t tot = 0;
  MGN_CUDA(cudaMemcpyAsync(&tot, offs + Ek, 8, cudaMemcpyDeviceToHost, stream));
  MGN_CUDA(cudaStreamSynchronize(stream));
  *total_host = tot + Ek;
  return MGN_OK;
}

// Phase 2: write the candidate keys [total] (then mgn_sort_unique_keys + mgn_keys_to_edge_index).
extern "C" int mgn_khop_expand(const int64_t* edge_index, int64_t E, const int64_t* edge_index_k, int64_t Ek,
                               int64_t num_nodes, const void* workspace, int64_t* keys, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  MGN_CHECK_ARG(edge_index && edge_index_k && workspace && keys, "khop_expand: null pointer");
  const int64_t* rowptr = reinterpret_cast<const int64_t*>(workspace);
  const int64_t* offs = rowptr + (num_nodes + 1) + (Ek + 1);
  khop_expand_kernel<<<blocks_for(Ek), TPB, 0, stream>>>(edge_index_k, Ek, edge_index, E, rowptr, offs, num_nodes, keys);
  MGN_LAUNCH_CHECK();
  return MGN_OK;
}

// ---- MeshMask filter_edges (meshmask.py:9-37) -----------------------------------
__global__ void fill_i64_kernel(int64_t* p, int64_t n, int64_t v) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) p[i] = v;
}
__global__ void scatter_iota_kernel(const int64_t* __restrict__ node_index, int64_t k, int64_t* __restrict__ map) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < k; i += (int64_t)gridDim.x * blockDim.x)
    map[node_index[i]] = i;
}
__global__ void filter_flag_kernel(const int64_t* __restrict__ ei, int64_t E, const int64_t* __restrict__ map,
                                   uint8_t* __restrict__ mask, int64_t* __restrict__ flag) {
  for (int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; e < E; e += (int64_t)gridDim.x * blockDim.x) {
    bool keep = map[ei[e]] >= 0 && map[ei[E + e]] >= 0;
    mask[e] = keep;
    flag[e] = keep;
  }
}
__global__ void filter_compact_kernel(const int64_t* __restrict__ ei, int64_t E, const int64_t* __restrict__ map,
                                      const int64_t* __restrict__ pos, int64_t Eout_cap,
                                      int64_t* __restrict__ out, int64_t* __restrict__ kept_ids) {
  for (int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; e < E; e += (int64_t)gridDim.x * blockDim.x) {
    if (pos[e + 1] != pos[e]) {
      int64_t o = pos[e];
      out[o] = map[ei[e]];
      out[Eout_cap + o] = map[ei[E + e]];
      kept_ids[o] = e;
    }
  }
}

extern "C" int64_t mgn_filter_workspace_bytes(int64_t E, int64_t num_nodes) {
  size_t a = 0;
  cub::DeviceScan::ExclusiveSum(nullptr, a, (const int64_t*)nullptr, (int64_t*)nullptr, E + 1);
  return (int64_t)((a + 255) / 256 * 256) + 8 * (num_nodes + 2 * (E + 1)) + 1024;
}

// out_edge_index is [2, E] capacity (row stride E); the caller narrows it to
// [:, :*count_host]; kept_ids [E] receives the original edge id of every kept edge
// (used to gather edge_attr rows); mask [E] uint8.  Original order is preserved.
extern "C" int mgn_filter_edges(const int64_t* edge_index, int64_t E, const int64_t* node_index, int64_t k,
                                int64_t num_nodes, int64_t* out_edge_index, int64_t* kept_ids, uint8_t* mask,
                                void* workspace, int64_t workspace_bytes, int64_t* count_host, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  MGN_CHECK_ARG(count_host && E >= 0 && k >= 0 && num_nodes >= 0, "filter_edges: bad sizes");
  if (E == 0) { *count_host = 0; return MGN_OK; }
  MGN_CHECK_ARG(edge_index && (k == 0 || node_index) && out_edge_index && kept_ids && mask && workspace,
                "filter_edges: null pointer");
  MGN_CHECK_ARG(workspace_bytes >= mgn_filter_workspace_bytes(E, num_nodes), "filter_edges: workspace too small");
  int64_t* map = reinterpret_cast<int64_t*>(workspace);
  int64_t* flag = map + num_nodes;
  int64_t* pos = flag + (E + 1);
  void* tmp = reinterpret_cast<void*>(((uintptr_t)(pos + (E + 1)) + 255) & ~(uintptr_t)255);
  size_t tmp_bytes = (size_t)workspace_bytes - ((uint8_t*)tmp - (uint8_t*)workspace);
  fill_i64_kernel<<<blocks_for(num_nodes), TPB, 0, stream>>>(map, num_nodes, -1);
  MGN_LAUNCH_CHECK();
  if (k > 0) {
    scatter_iota_kernel<<<blocks_for(k), TPB, 0, stream>>>(node_index, k, map);
    MGN_LAUNCH_CHECK();
  }
  MGN_CUDA(cudaMemsetAsync(flag + E, 0, 8, stream));
  filter_flag_kernel<<<blocks_for(E), TPB, 0, stream>>>(edge_index, E, map, mask, flag);
  MGN_LAUNCH_CHECK();
  MGN_CUDA(cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, flag, pos, (int)(E + 1), stream));
  filter_compact_kernel<<<blocks_for(E), TPB, 0, stream>>>(edge_index, E, map, pos, E, out_edge_index, kept_ids);
  MGN_LAUNCH_CHECK();
  MGN_CUDA(cudaMemcpyAsync(count_host, pos + E, 8, cudaMemcpyDeviceToHost, stream));
  MGN_CUDA(cudaStreamSynchronize(stream));
  return MGN_OK;
}

// ---- edge features: [pos[row]-pos[col], ||pos[col]-pos[row]||_2], D = 2 or 3 -----
__global__ void edge_features_kernel(const float* __restrict__ pos, int D, const int64_t* __restrict__ ei,
                                     int64_t E, float* __restrict__ out) {
  for (int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; e < E; e += (int64_t)gridDim.x * blockDim.x) {
    int64_t r = ei[e], c = ei[E + e];
    float ss = 0.f;
    for (int d = 0; d < D; ++d) {
      float pr = pos[r * D + d], pc = pos[c * D + d];
      out[e * (D + 1) + d] = __fsub_rn(pr, pc);
      float df = __fsub_rn(pc, pr);
      ss = __fadd_rn(ss, __fmul_rn(df, df));
    }
    out[e * (D + 1) + D] = __fsqrt_rn(ss);
  }
}

extern "C" int mgn_edge_features(const float* pos, int dim, const int64_t* edge_index, int64_t E, float* out,
                                 void* stream) {
  MGN_CHECK_ARG(dim >= 1 && dim <= 3 && E >= 0, "edge_features: bad sizes");
  if (E == 0) return MGN_OK;
  MGN_CHECK_ARG(pos && edge_index && out, "edge_features: null pointer");
  edge_features_kernel<<<blocks_for(E), TPB, 0, (cudaStream_t)stream>>>(pos, dim, edge_index, E, out);
  MGN_LAUNCH_CHECK();
  return MGN_OK;
}
