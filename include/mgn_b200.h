/* mgn_b200.h -- C ABI of libmgn_b200.so: the B200 (sm_100a) MeshGraphNet processor path.
 *
 * The reference (cviviers/graph-physics) is pure Python; it has no FFI.  The boundary
 * it offers for this path is the nn.Module API of graphphysics.models
 * (GraphNetBlock.forward, graphphysics/models/layers.py:667-701;
 * EncodeProcessDecode.forward, graphphysics/models/processors.py:111-137) and the
 * set-up functions of graphphysics/utils and graphphysics/dataset.  The Python host in
 * graph_physics_b200/ mirrors those signatures and binds the entry points below with
 * ctypes (INTEGRATION.md shows the binding).  Each entry point cites the reference
 * code it replaces.
 *
 * Conventions
 *  - plain pointers and sizes only; all pointers are DEVICE pointers unless named *_host
 *  - the library never allocates or frees device memory: outputs and workspaces are
 *    provided by the caller (sizes from the *_workspace_bytes queries)
 *  - `stream` is a cudaStream_t passed as void*; nothing synchronises unless stated
 *  - return value: 0 ok; < 0 bad argument / unsupported shape; > 0 a cudaError_t.
 *    mgn_last_error() returns the thread-local message.  Nothing throws or exits.
 *  - edge_index is int64 [2, E] row-major: row 0 = sender, row 1 = receiver.
 */
#ifndef MGN_B200_H
#define MGN_B200_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

const char* mgn_last_error(void);
int mgn_version(void);
int mgn_check_device(void); /* 0 iff the current device is sm_100 */

/* ------------------------------------------------------------------ graph set-up */
/* meshdata_to_graph tet->4 triangles, graphphysics/utils/torch_graph.py:170-186.
 * cells [4,T] int64 -> face [3,4T] int64. */
int mgn_tets_to_faces(const int64_t* cells, int64_t T, int64_t* face, void* stream);
/* T.FaceToEdge slices + reversed pairs (preprocessing.py:410,429): face [3,F] -> 6F keys
 * row*N+col. */
int mgn_face_keys(const int64_t* face, int64_t F, int64_t num_nodes, int64_t* keys, void* stream);
/* to_undirected / coalesce first half (preprocessing.py:137): edge_index -> E (or 2E) keys. */
int mgn_edge_keys(const int64_t* edge_index, int64_t E, int64_t num_nodes, int both_directions,
                  int64_t* keys, void* stream);
/* coalesce: sort ascending + unique.  keys_in is clobbered.  Synchronises the stream to
 * return the distinct count in *count_host. */
int64_t mgn_sort_unique_workspace_bytes(int64_t n);
int mgn_sort_unique_keys(int64_t* keys_in, int64_t n, int64_t num_nodes, int64_t* keys_out,
                         void* workspace, int64_t workspace_bytes, int64_t* count_host, void* stream);
int mgn_keys_to_edge_index(const int64_t* keys, int64_t E, int64_t num_nodes, int64_t* edge_index,
                           void* stream);
/* receiver-sorted CSR (stable) replacing the scatter-sum at edge_index[1] of
 * MessagePassing.propagate (layers.py:694-696).  rowptr [N+1], perm/senders/receivers [E], int32. */
int64_t mgn_csr_workspace_bytes(int64_t E);
int mgn_build_csr(const int64_t* edge_index, int64_t E, int64_t num_nodes, int32_t* rowptr, int32_t* perm,
                  int32_t* senders, int32_t* receivers, void* workspace, int64_t workspace_bytes,
                  void* stream);
/* compute_k_hop_edge_index, graphphysics/utils/torch_graph.py:16-53: one hop step =
 * mgn_khop_count (synchronises, returns candidate count) + mgn_khop_expand + sort_unique. */
int64_t mgn_khop_workspace_bytes(int64_t E, int64_t Ek, int64_t num_nodes);
int mgn_khop_count(const int64_t* edge_index, int64_t E, const int64_t* edge_index_k, int64_t Ek,
                   int64_t num_nodes, void* workspace, int64_t workspace_bytes, int64_t* total_host,
                   void* stream);
int mgn_khop_expand(const int64_t* edge_index, int64_t E, const int64_t* edge_index_k, int64_t Ek,
                    int64_t num_nodes, const void* workspace, int64_t* keys, void* stream);
/* filter_edges, graphphysics/utils/meshmask.py:9-37.  out_edge_index has row stride E. */
int64_t mgn_filter_workspace_bytes(int64_t E, int64_t num_nodes);
int mgn_filter_edges(const int64_t* edge_index, int64_t E, const int64_t* node_index, int64_t k,
                     int64_t num_nodes, int64_t* out_edge_index, int64_t* kept_ids, uint8_t* mask,
                     void* workspace, int64_t workspace_bytes, int64_t* count_host, void* stream);
/* add_edge_features = T.Cartesian(norm=False) + T.Distance(norm=False),
 * graphphysics/dataset/preprocessing.py:16-23.  pos [N,dim] f32 -> out [E,dim+1] f32. */
int mgn_edge_features(const float* pos, int dim, const int64_t* edge_index, int64_t E, float* out,
                      void* stream);

/* ------------------------------------------------------------------ self test */
/* one-tile tcgen05 product pinning the shared-memory descriptor conventions (tests only) */
int mgn_selftest_umma(const void* a, const void* b, float* d, int a_cols, int b_cols, int a_mn, int b_mn,
                      int n, int ksteps, int b_col0, int use_ones, int repeat, int lbo_a, int sbo_a,
                      int lbo_b, int sbo_b, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* MGN_B200_H */
