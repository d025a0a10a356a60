// graph_setup.cu -- one-time integer graph set-up on the device:
// faces -> directed edge keys -> radix sort -> unique -> PyG-ordered edge_index,
// receiver-sorted CSR, K-hop expansion, MeshMask edge filtering, Cartesian/Distance
// edge features.  All index results are bit-exact with the reference's
// torch_geometric / ATen path (ascending (row, col) order, int64 edge_index).
//
// Replaces: T.FaceToEdge + to_undirected (graphphysics/dataset/preprocessing.py:137,410,429),
// compute_k_hop_edge_index (graphphysics/utils/torch_graph.py:16-53),
// filter_edges (graphphysics/utils/meshmask.py:9-37), add_edge_features
// (preprocessing.py:16-23).  HBM-bound integer work: coalesced streams + CUB radix sort.
#include <cub/cub.cuh>

#include "common.cuh"

static constexpr int TPB = 256;
static inline int blocks_for(int64_t n) { return (int)std::min<int64_t>(ceil_div64(n, TPB), 148 * 16); }

// ---- S1: tets -> 4 triangles, block-wise (torch_graph.py:170-186)
__global__ void tets_to_faces_kernel(const int64_t* __restrict__ cells /*[4,T]*/, int64_t T,
                                     int64_t* __restrict__ face /*[3,4T]*/) {
  const int64_t F = 4 * T;
  for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < T; t += (int64_t)gridDim.x * blockDim.x) {
    int64_t c0 = cells[t], c1 = cells[T + t], c2 = cells[2 * T + t], c3 = cells[3 * T + t];
    face[t] = c0;            face[F + t] = c1;            face[2 * F + t] = c2;
    face[T + t] = c1;        face[F + T + t] = c2;        face[2 * F + T + t] = c3;
    face[2 * T + t] = c2;    face[F + 2 * T + t] = c3;    face[2 * F + 2 * T + t] = c0;
    face[3 * T + t] = c3;    face[F + 3 * T + t] = c0;    face[2 * F + 3 * T + t] = c1;
  }
}

// ---- S2: face [3,F] -> 6F keys row*N+col (both directions of the 3 sides)
__global__ void face_keys_kernel(const int64_t* __restrict__ face, int64_t F, int64_t N,
                                 int64_t* __restrict__ keys) {
  for (int64_t f = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; f < F; f += (int64_t)gridDim.x * blockDim.x) {
    int64_t a = face[f], b = face[F + f], c = face[2 * F + f];
    keys[f] = a * N + b;          keys[F + f] = b * N + c;      keys[2 * F + f] = a * N + c;
    keys[3 * F + f] = b * N + a;  keys[4 * F + f] = c * N + b;  keys[5 * F + f] = c * N + a;
  }
}

// edge_index [2,E] -> keys, optionally appending the reversed pairs (to_undirected)
__global__ void edge_keys_kernel(const int64_t* __restrict__ ei, int64_t E, int64_t N, int both,
                                 int64_t* __restrict__ keys) {
  for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < E; k += (int64_t)gridDim.x * blockDim.x) {
    int64_t r = ei[k], c = ei[E + k];
    keys[k] = r * N + c;
    if (both) keys[E + k] = c * N + r;
  }
}

__global__ void keys_to_edges_kernel(const int64_t* __restrict__ keys, int64_t E, int64_t N,
                                     int64_t* __restrict__ ei) {
  for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < E; k += (int64_t)gridDim.x * blockDim.x) {
    int64_t key = keys[k];
    ei[k] = key / N;
    ei[E + k] = key % N;
  }
}

static int key_bits(int64_t N) {
  int bits = 1;
  // keys < N*N
  unsigned __int128 lim = (unsigned __int128)N * (unsigned __int128)N;
  while (bits < 64 && (((unsigned __int128)1) << bits) < lim) ++bits;
  return bits;
}

extern "C" int64_t mgn_sort_unique_workspace_bytes(int64_t n) {
  size_t a = 0, b = 0;
  cub::DeviceRadixSort::SortKeys(nullptr, a, (const int64_t*)nullptr, (int64_t*)nullptr, n);
  cub::DeviceSelect::Unique(nullptr, b, (const int64_t*)nullptr, (int64_t*)nullptr, (int64_t*)nullptr, n);
  size_t w = a > b ? a : b;
  return (int64_t)((w + 255) / 256 * 256 + 256);
}

// keys_in [n] (clobbered as scratch), keys_out [n]; *count_host receives the number
// of distinct keys (the stream is synchronised to read it).  Sorted ascending.
extern "C" int mgn_sort_unique_keys(int64_t* keys_in, int64_t n, int64_t num_nodes, int64_t* keys_out,
                                    void* workspace, int64_t workspace_bytes, int64_t* count_host,
                                    void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  MGN_CHECK_ARG(n >= 0 && num_nodes > 0 && count_host, "sort_unique: bad sizes");
  if (n == 0) { *count_host = 0; return MGN_OK; }
  MGN_CHECK_ARG(keys_in && keys_out && workspace, "sort_unique: null pointer");
  MGN_CHECK_ARG(n < (int64_t)1 << 31, "sort_unique: n too large");
  MGN_CHECK_ARG(workspace_bytes >= mgn_sort_unique_workspace_bytes(n), "sort_unique: workspace too small");
  int64_t* d_count = reinterpret_cast<int64_t*>(workspace);
  void* tmp = reinterpret_cast<uint8_t*>(workspace) + 256;
  size_t tmp_bytes = (size_t)workspace_bytes - 256;
  MGN_CUDA(cub::DeviceRadixSort::SortKeys(tmp, tmp_bytes, keys_in, keys_out, (int)n, 0, key_bits(num_nodes), stream));
  MGN_CUDA(cub::DeviceSelect::Unique(tmp, tmp_bytes, keys_out, keys_in, d_count, (int)n, stream));
  MGN_CUDA(cudaMemcpyAsync(keys_out, keys_in, sizeof(int64_t) * n, cudaMemcpyDeviceToDevice, stream));
  MGN_CUDA(cudaMemcpyAsync(count_host, d_count, sizeof(int64_t), cudaMemcpyDeviceToHost, stream));
  MGN_CUDA(cudaStreamSynchronize(stream));
  return MGN_OK;
}

extern "C" int mgn_tets_to_faces(const int64_t* cells, int64_t T, int64_t* face, void* stream) {
  MGN_CHECK_ARG(T >= 0 && (T == 0 || (cells && face)), "tets_to_faces: bad args");
  if (T == 0) return MGN_OK;
  tets_to_faces_kernel<<<blocks_for(T), TPB, 0, (cudaStream_t)stream>>>(cells, T, face);
  MGN_LAUNCH_CHECK();
  return MGN_OK;
}

extern "C" int mgn_face_keys(const int64_t* face, int64_t F, int64_t num_nodes, int64_t* keys, void* stream) {
  MGN_CHECK_ARG(F >= 0 && num_nodes > 0 && (F == 0 || (face && keys)), "face_keys: bad args");
  if (F == 0) return MGN_OK;
  face_keys_kernel<<<blocks_for(F), TPB, 0, (cudaStream_t)stream>>>(face, F, num_nodes, keys);
  MGN_LAUNCH_CHECK();
  return MGN_OK;
}

extern "C" int mgn_edge_keys(const int64_t* edge_index, int64_t E, int64_t num_nodes, int both_directions,
                             int64_t* keys, void* stream) {
  MGN_CHECK_ARG(E >= 0 && num_nodes > 0 && (E == 0 || (edge_index && keys)), "edge_keys: bad args");
  if (E == 0) return MGN_OK;
  edge_keys_kernel<<<blocks_for(E), TPB, 0, (cudaStream_t)stream>>>(edge_index, E, num_nodes, both_directions, keys);
  MGN_LAUNCH_CHECK();
  return MGN_OK;
}

extern "C" int mgn_keys_to_edge_index(const int64_t* keys, int64_t E, int64_t num_nodes, int64_t* edge_index,
                                      void* stream) {
  MGN_CHECK_ARG(E >= 0 && num_nodes > 0 && (E == 0 || (keys && edge_index)), "keys_to_edge_index: bad args");
  if (E == 0) return MGN_OK;
  keys_to_edges_kernel<<<blocks_for(E), TPB, 0, (cudaStream_t)stream>>>(keys, E, num_nodes, edge_index);
  MGN_LAUNCH_CHECK();
  return MGN_OK;
}

// ---- receiver-sorted CSR -------------------------------------------------------
__global__ void csr_prepare_kernel(const int64_t* __restrict__ ei, int64_t E, int32_t* __restrict__ recv,
                                   int32_t* __restrict__ iota) {
  for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < E; k += (int64_t)gridDim.x * blockDim.x) {
    recv[k] = (int32_t)ei[E + k];
    iota[k] = (int32_t)k;
  }
}
__global__ void csr_finish_kernel(const int64_t* __restrict__ ei, int64_t E, int64_t N,
                                  const int32_t* __restrict__ recv_sorted, const int32_t* __restrict__ perm,
                                  int32_t* __restrict__ senders, int32_t* __restrict__ rowptr) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < E; k += stride)
    senders[k] = (int32_t)ei[perm[k]];
  for (int64_t n = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; n <= N; n += stride) {
    int64_t lo = 0, hi = E;  // lower_bound(recv_sorted, n)
    while (lo < hi) {
      int64_t mid = (lo + hi) >> 1;
      if (recv_sorted[mid] < n) lo = mid + 1; else hi = mid;
    }
    rowptr[n] = (int32_t)lo;
  }
}

extern "C" int64_t mgn_csr_workspace_bytes(int64_t E) {
  size_t a = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, a, (const int32_t*)nullptr, (int32_t*)nullptr,
                                  (const int32_t*)nullptr, (int32_t*)nullptr, E);
  return (int64_t)((a + 255) / 256 * 256) + 2 * ((E * 4 + 255) / 256 * 256) + 256;
}

// Stable sort of the caller's edges by receiver (edge_index[1]).
//   rowptr [N+1], perm [E] (perm[k] = caller edge id of CSR edge k), senders [E],
//   receivers [E] (both in CSR order); all int32.
extern "C" int mgn_build_csr(const int64_t* edge_index, int64_t E, int64_t num_nodes, int32_t* rowptr,
                             int32_t* perm, int32_t* senders, int32_t* receivers, void* workspace,
                             int64_t workspace_bytes, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  MGN_CHECK_ARG(E >= 0 && num_nodes > 0 && rowptr, "build_csr: bad sizes");
  MGN_CHECK_ARG(E < ((int64_t)1 << 31) && num_nodes < ((int64_t)1 << 31), "build_csr: int32 overflow");
  if (E == 0) {
    MGN_CUDA(cudaMemsetAsync(rowptr, 0, sizeof(int32_t) * (num_nodes + 1), stream));
    return MGN_OK;
  }
  MGN_CHECK_ARG(edge_index && perm && senders && receivers && workspace, "build_csr: null pointer");
  MGN_CHECK_ARG(workspace_bytes >= mgn_csr_workspace_bytes(E), "build_csr: workspace too small");
  const int64_t seg = (E * 4 + 255) / 256 * 256;
  int32_t* recv_in = reinterpret_cast<int32_t*>(workspace);
  int32_t* iota = reinterpret_cast<int32_t*>(reinterpret_cast<uint8_t*>(workspace) + seg);
  void* tmp = reinterpret_cast<uint8_t*>(workspace) + 2 * seg;
  size_t tmp_bytes = (size_t)(workspace_bytes - 2 * seg);
  csr_prepare_kernel<<<blocks_for(E), TPB, 0, stream>>>(edge_index, E, recv_in, iota);
  MGN_LAUNCH_CHECK();
  int bits = 1;
  while (bits < 31 && ((int64_t)1 << bits) < num_nodes) ++bits;
  MGN_CUDA(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, recv_in, receivers, iota, perm, (int)E, 0, bits, stream));
  csr_finish_kernel<<<blocks_for(std::max<int64_t>(E, num_nodes + 1)), TPB, 0, stream>>>(
      edge_index, E, num_nodes, receivers, perm, senders, rowptr);
  MGN_LAUNCH_CHECK();
  return MGN_OK;
}
