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
int mgn_version(void);      /* 200 = 0.2.0 (round 2: mgn_edge_fwd removed; carry / rowptr / norm_dim arguments; chained entry points) */
int mgn_check_device(void); /* 0 iff the current device is sm_100 */
/* Every mbarrier wait inside the kernels is bounded (4 s): a wait that starves writes {source line, block, thread,
 * barrier, parity} to host-mapped memory and traps, so the launch fails with a cudaError instead of spinning.
 * mgn_hang_report copies a one-line description of that record into buf (returns 1 if there is one, else 0). */
int mgn_hang_report(char* buf, int n);

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

/* add_world_edges, graphphysics/dataset/preprocessing.py:92-140: the radius search that replaces
 * scipy.spatial.cKDTree.query_pairs + the OBSTACLE/NORMAL masks.  Uniform cell list (cell edge 1.001 radius):
 *   1. mgn_world_cell_keys -> keys [N] int64 = ((sample*DX + cx)*DY + cy)*DZ + cz; the caller sorts them
 *      (keys_sorted ascending, order [N] int32 = node at each sorted position);
 *   2. mgn_world_pairs with offs == NULL -> counts [N] int64 (pairs found per first-set node);
 *   3. mgn_world_pairs with offs = exclusive sum of counts -> pairs [2,M] int64 (row 0 = first-set node).
 * pos [N,dim] f32 (dim 2 or 3); cls [N] u8: 1 = first set (OBSTACLE), 2 = second set (NORMAL), other = ignored;
 * sample [N] int32 or NULL: pairs only join nodes of the same sample; origin[3], dims[3] are HOST arrays
 * (per-axis minimum; floor((max-origin)/(1.001 radius)) + 1).  Test: squared distance of the float64-converted
 * coordinates <= radius^2 (scipy's).  The caller then runs mgn_edge_keys/mgn_sort_unique_keys (to_undirected) on
 * pairs ++ mesh edges exactly as the reference does. */
int mgn_world_cell_keys(const float* pos, int dim, int64_t N, const int32_t* sample, const double* origin,
                        double radius, const int64_t* dims, int64_t* keys, void* stream);
int mgn_world_pairs(const float* pos, int dim, int64_t N, const uint8_t* cls, const int32_t* sample,
                    const int64_t* keys_sorted, const int32_t* order, const double* origin, double radius,
                    const int64_t* dims, int64_t* counts, const int64_t* offs, int64_t* pairs, int64_t M,
                    void* stream);

/* ------------------------------------------------------------------ message-passing step
 * One GraphNetBlock (graphphysics/models/layers.py:622-746) with hidden_size 128, nb_of_layers 4,
 * layer_norm (RMSNorm, layers.py:49-74).  Latents are fp32 [rows,128] row-major.  Products run on
 * tcgen05 with fp16 operands and fp32 accumulation; residual streams, RMSNorm and the aggregation
 * are fp32.  Edge rows are in receiver-sorted CSR order (mgn_build_csr), or in the caller's order
 * when `rowmap` (= perm) is given.  The first edge layer is split: the caller runs mgn_p_fwd per
 * node, the edge kernel adds P[:, :128][receiver] + P[:, 128:][sender] to e We^T.
 */
/* bytes of one block's packed fp16 weight tiles (edge We,W2,W3,W4 | [W_i;W_j] | node W1',W2',W3',W4') */
int64_t mgn_packed_block_bytes(void);
/* build_mlp weights (layers.py:137-183; fp32, [out,in] row-major: eW1 [128,384], nW1 [128,256],
 * others [128,128]) -> 128-byte-swizzled fp16 panel tiles in `packed`. */
int mgn_pack_block(const float* eW1, const float* eW2, const float* eW3, const float* eW4,
                   const float* nW1, const float* nW2, const float* nW3, const float* nW4,
                   void* packed, void* stream);
/* the same for K blocks in one launch: weights_host = HOST array of K * 8 DEVICE pointers (per block the eight
 * arguments of mgn_pack_block in order); packed = [K][mgn_packed_block_bytes()] */
int mgn_pack_blocks(int K, const float* const* weights_host, void* packed, void* stream);
/* P [N,256] fp16 = [x W_i^T + b1 | x W_j^T]: the x_i / x_j columns of edge_block layer 1
 * (cat order [edge_attr, x_i, x_j], layers.py:703-719). */
int mgn_p_fwd(int64_t N, const float* x, const void* packed, const float* b1_edge, void* P, void* stream);
/* edge update + sum aggregation at the receiver (layers.py:687-698): e_out = e_in + edge_block(...),
 * aggr [N,128] = segmented sum of the PRE-residual messages (zeroed here).  save_h* (optional,
 * [E,128] fp16) receive the hidden activations.  The fp16 operand-tile images (ceil(E/128) * 32768 bytes,
 * CSR tile order) tie consecutive steps and the backward together: e16_in (optional) = images of e_in left
 * by the previous step (operand tiles then arrive by bulk copy; e_in is read for the residual only);
 * e16_out (optional, used when e16_in is null) receives the images of e_in; e16_next (optional) receives
 * the images of e_out.  mgn_edge_bwd recomputes from the images of e_in.  out_in_csr_order != 0 with a rowmap:
 * e_in is gathered through the rowmap but e_out is written in CSR order (the entry permutation of a processor
 * loop fused into its first step). */
/* e16lo_next: tile images of  e_out - fp16(e_out)  (fp16), i.e. together with e16_next the
 * image + remainder form of the residual stream that mgn_edge_fwd_hl consumes.  With both images requested e_out
 * may be NULL (the fp32 rows are then not written at all).
 * carry: fp32 workspace [ceil(E/8),128].  The sum at the receivers (PyG scatter-sum, layers.py:694-696) is
 * DETERMINISTIC and atomics-free: the receiver-sorted edge list is cut into 8-row slots, a run of equal receivers
 * that starts in a slot is stored to its node's row of aggr, the partial sums of its continuation in later slots go
 * to carry[slot] and a second kernel adds them in slot order (one writer per node).
 * rowptr: [N+1] receiver CSR or NULL: with it the second kernel also zeroes the rows of nodes without in-edges and
 * aggr needs no initialisation (NULL: aggr is cleared with a memset first).
 * norm_dim: the model's true hidden size d for RMSNorm's d^-1/2 (layers.py:66); 128 unless the caller zero-pads a
 * narrower model into the 128-wide kernels. */
int mgn_edge_fwd2(int64_t E, int64_t N, const float* e_in, const int32_t* rowmap, float* e_out,
                  const void* P, const int32_t* senders, const int32_t* receivers, float* aggr,
                  const void* packed, const float* b2, const float* b3, const float* b4,
                  const float* scale, void* save_h1, void* save_h2, void* save_h3, void* e16_out,
                  const void* e16_in, void* e16_next, void* e16lo_next, int out_in_csr_order, float* carry,
                  const int32_t* rowptr, int norm_dim, void* stream);
/* The same edge update (layers.py:687-698) for steps 1..K-1 with the residual stream held as two fp16 tile-image
 * streams, hi = fp16(e) (= the operand images mgn_edge_bwd recomputes from) and lo = fp16(e - hi): e = hi + lo
 * carries >= 21 significant bits of the fp32 stream, the residual add is done in fp32, and the stream moves only by
 * 32 KB bulk copies (csrc/edge_fwd_hl.cu).  hi_out / lo_out (both or neither) receive the images of e_out; e_out
 * (fp32 rows, CSR order) is optional -- only a caller that wants the edge latents back needs it.
 * carry: [ceil(E/32),128] fp32 workspace, norm_dim: as for mgn_edge_fwd2 (slots of 32 rows here).
 * h1_save, h2_save, h3_save (all three or all NULL): tile images (ceil(E/128) * 32768 bytes each) receiving the edge
 * MLP's hidden activations for mgn_edge_bwd_saved. */
int mgn_edge_fwd_hl(int64_t E, int64_t N, const void* hi_in, const void* lo_in, void* hi_out, void* lo_out,
                    float* e_out, const void* P, const int32_t* senders, const int32_t* receivers, float* aggr,
                    const void* packed, const float* b2, const float* b3, const float* b4, const float* scale,
                    float* carry, const int32_t* rowptr, int norm_dim, void* h1_save, void* h2_save, void* h3_save,
                    void* stream);
/* mgn_edge_fwd_hl without a caller-provided workspace and WITHOUT ANY ALLOCATION (no cudaMalloc / cudaFree / allocating
 * routine), for callers that cannot allocate: code under CUDA-graph capture, tight serving loops, harnesses that forbid
 * it.  No counterpart in the reference.  The carry scratch is a file-scope
 * device array of the library: E <= mgn_noalloc_max_edges() = 1,048,576 edges (16 MiB, held by every GPU that loads the
 * library until exit); larger E returns -3 with a message.  Every call shares that array (contents of the last call
 * remain; none of it needs clearing): calls that overlap on the GPU race -- serialise them.  Same results as
 * mgn_edge_fwd_hl. */
int64_t mgn_noalloc_max_edges(void);
int mgn_edge_fwd_hl_noalloc(int64_t E, int64_t N, const void* hi_in, const void* lo_in, void* hi_out, void* lo_out,
                            float* e_out, const void* P, const int32_t* senders, const int32_t* receivers,
                            float* aggr, const void* packed, const float* b2, const float* b3, const float* b4,
                            const float* scale, const int32_t* rowptr, int norm_dim, void* stream);
/* node update (layers.py:733-746, 699): x_out = x + node_block(cat[x, aggr]).  save_h* (optional, all three or
 * none) receive the hidden activations mgn_node_bwd needs as fp16 TILE IMAGES: ceil(N/128) * 32768 bytes each, one
 * 128-byte-swizzled 128 x 128 operand tile per 128 rows, bulk-stored as it sits in shared memory. */
int mgn_node_fwd(int64_t N, const float* x, const float* aggr, float* x_out, const void* packed,
                 const float* b1, const float* b2, const float* b3, const float* b4,
                 const float* scale, void* save_h1, void* save_h2, void* save_h3, void* stream);
/* mgn_node_fwd with flags: bit 0 = the block has no RMSNorm (build_mlp(..., layer_norm=False), the decoder of
 * EncodeProcessDecode, processors.py:85-90): x_out = x + MLP(cat[x, aggr]); `scale` is ignored. */
int mgn_node_fwd2(int64_t N, const float* x, const float* aggr, float* x_out, const void* packed,
                  const float* b1, const float* b2, const float* b3, const float* b4,
                  const float* scale, void* save_h1, void* save_h2, void* save_h3, int flags, int norm_dim,
                  void* stream);

/* ------------------------------------------------------------------ backward of the step
 * What torch.autograd derives from GraphNetBlock.forward in the reference.  Parameter gradients of
 * one block accumulate (red.add) into a flat fp32 buffer of mgn_grad_block_floats() floats laid out
 * in the block's state-dict order (edge_block.0.weight, .0.bias, ..., .7.scale, node_block...);
 * the caller zeroes it.  Upstream gradients may be pre-scaled by any power of two. */
int64_t mgn_grad_block_floats(void);
/* node_block backward from the fp16 hidden saves (the tile images mgn_node_fwd wrote): g_x = g_xout + dz1 W1'[:, :128],
 * g_aggr = dz1 W1'[:, 128:], dz1 [N,128] fp16; dW2'..dW4', all biases, dscale. */
int mgn_node_bwd(int64_t N, const float* g_xout, const void* h1, const void* h2, const void* h3,
                 float* g_x, float* g_aggr, void* dz1, const void* packed, const float* b4,
                 const float* scale, float* grad_block, void* stream);
/* mgn_node_bwd with flags: bit 0 = no RMSNorm (see mgn_node_fwd2): dz4 = g_xout, no scale gradient. */
int mgn_node_bwd2(int64_t N, const float* g_xout, const void* h1, const void* h2, const void* h3,
                  float* g_x, float* g_aggr, void* dz1, const void* packed, const float* b4,
                  const float* scale, float* grad_block, int flags, int norm_dim, void* stream);
/* Tile-level chaining of two launches (no counterpart in the reference: it is what removes the wave tail of the
 * persistent node kernels -- 481 tiles on 148 SMs).  mgn_node_bwd3 = mgn_node_bwd2 that also publishes
 * ready_out[tile] = 1 (int32, one per 128-row tile) after the tile's dz1 / g_x / g_aggr rows are written;
 * mgn_node_wgrad1c = mgn_node_wgrad1 that, launched DIRECTLY after it on the same stream, takes its tiles from
 * sched[0] in index order, waits for ready_in[tile] (bounded wait, mgn_hang_report) and clears it, and only then waits
 * for the whole previous launch -- its CTAs start on the SMs node_bwd's last wave leaves idle.  ready / sched
 * ({next tile, CTAs done}) start at zero and are left at zero.  NULL for both = plain stream order. */
int mgn_node_bwd3(int64_t N, const float* g_xout, const void* h1, const void* h2, const void* h3,
                  float* g_x, float* g_aggr, void* dz1, const void* packed, const float* b4,
                  const float* scale, float* grad_block, int flags, int norm_dim, int32_t* ready_in, int32_t* ready_out,
                  void* stream);
int mgn_node_wgrad1c(int64_t N, const void* dz1, const float* x, const float* aggr, float* grad_block,
                     int32_t* ready_in, int32_t* sched, void* stream);
/* The same pair in the forward: mgn_node_fwd3 publishes ready_out[tile] after the tile's x_out rows, mgn_p_fwd_c (the
 * partial products of the NEXT step, launched after it on the same stream) consumes them. */
int mgn_node_fwd3(int64_t N, const float* x, const float* aggr, float* x_out, const void* packed,
                  const float* b1, const float* b2, const float* b3, const float* b4,
                  const float* scale, void* save_h1, void* save_h2, void* save_h3, int flags, int norm_dim,
                  int32_t* ready_out, void* stream);
int mgn_p_fwd_c(int64_t N, const float* x, const void* packed, const float* b1_edge, void* P, int32_t* ready_in,
                int32_t* sched, void* stream);
/* And p_bwd(k) -> node_bwd(k-1): mgn_p_bwd_c publishes ready_out[tile] after the tile's g_x rows; mgn_node_bwd3's
 * ready_in (same buffer) makes it wait per tile instead of for the whole p_bwd launch.  Both kernels give tile t to CTA
 * t mod grid, so this pair needs no tile counter. */
int mgn_p_bwd_c(int64_t N, const void* dP, const float* x, float* g_x, const void* packed, float* grad_block,
                int32_t* ready_out, void* stream);
/* dW1' += dz1^T [x | aggr] */
int mgn_node_wgrad1(int64_t N, const void* dz1, const float* x, const float* aggr, float* grad_block,
                    void* stream);
/* edge_block + aggregation backward; recomputes the edge MLP from e16 (the forward's tile images)
 * or, if e16 is null, from the fp32 rows e_in (no hidden-activation saves either way).
 * g = g_eout + g_aggr[receiver]; g_e = g_eout + dz1 We (g_e may be the same buffer as g_eout; g_eout
 * may be null = no upstream edge gradient).  dz16_out (ceil(E/128) * 32768 bytes) receives dz1, the
 * gradient at the first layer's pre-activation, as fp16 tile images in CSR order: the backward of the
 * gathers x[col], x[row] (layers.py:687-691) is finished by mgn_p_bwd without atomics. */
int mgn_edge_bwd(int64_t E, int64_t N, const float* e_in, const void* e16, const int32_t* rowmap,
                 const float* g_eout, const float* g_aggr, const void* P, const int32_t* senders,
                 const int32_t* receivers, float* g_e, void* dz16_out, const void* packed, const float* b2,
                 const float* b3, const float* b4, const float* scale, float* grad_block, int norm_dim,
                 void* stream);
/* mgn_edge_bwd for the processor loop when mgn_edge_fwd_hl kept the edge MLP's hidden activations (h1_save ..
 * h3_save): the autograd of layers.py:703-719, 694-696 WITHOUT recomputing layers 1-3.  g_e [E,128] (CSR order) holds
 * the upstream edge gradient on entry (zeros if none) and is updated in place to g_e + dz1 We. */
int mgn_edge_bwd_saved(int64_t E, int64_t N, const void* h1, const void* h2, const void* h3, const float* g_aggr,
                       const int32_t* receivers, float* g_e, void* dz16_out, const void* packed, const float* b4,
                       const float* scale, float* grad_block, int norm_dim, void* stream);
/* edge_block layer-1 weight gradient, edge-latent columns (edge_block.0.weight[:, 0:128]):
 * dWe += dz1^T e over all edges, streamed from the tile images of dz1 (mgn_edge_bwd) and of e (mgn_edge_fwd). */
int mgn_edge_wgrad1(int64_t E, const void* dz16, const void* e16, float* grad_block, void* stream);
/* backward of the two gathers: dP [N,256] fp16; dP[n, :128] = sum of dz1 over CSR positions
 * rowptr[n]..rowptr[n+1] (edges n receives), dP[n, 128:] = sum over out_pos[rowptr_s[n]..rowptr_s[n+1]]
 * (edges n sends; lists = mgn_build_csr of the flipped CSR-ordered pairs).  fp32 sums in a fixed
 * order: bitwise reproducible, no atomics. */
int mgn_dp_gather(int64_t N, const void* dz16, const int32_t* rowptr, const int32_t* rowptr_s,
                  const int32_t* out_pos, void* dP, void* stream);
/* backward of mgn_p_fwd: g_x += dP [W_i ; W_j]; dW_i, dW_j, db1 into grad_block */
int mgn_p_bwd(int64_t N, const void* dP, const float* x, float* g_x, const void* packed,
              float* grad_block, void* stream);

/* ------------------------------------------------------------------ developer hooks */
/* per-phase clock64 stamps of CTA 0 of the next edge/node kernels into `buf` (device, `cap`
 * entries; null disables) -- tools/timeline.py */
int mgn_debug_set_timeline(long long* buf, int cap);

/* ------------------------------------------------------------------ self test */
/* one-tile tcgen05 product pinning the shared-memory descriptor conventions (tests only) */
int mgn_selftest_umma(const void* a, const void* b, float* d, int a_cols, int b_cols, int a_mn, int b_mn,
                      int n, int ksteps, int b_col0, int use_ones, int repeat, int lbo_a, int sbo_a,
                      int lbo_b, int sbo_b, void* stream);
/* probe of shapes the fused kernels do not use yet: UMMA M = 64 (m), A operand from tensor memory (a_tmem 1/2 =
 * packing of two fp16 per 32-bit column); d [128 TMEM lanes][n] (tools/umma_probe2.py) */
int mgn_selftest_umma2(const void* a, const void* b, float* d, int n, int ksteps, int m, int a_tmem, void* stream);
/* probe of the CTA-pair product planned for the round-2 edge kernels: D[256][n] = A[256][16 ksteps] * B[n][..]^T by one
 * tcgen05.mma.cta_group::2 stream on a cluster of two CTAs, each holding half of B (a_tmem 1 = A operand in tensor
 * memory).  Not used by any product kernel; run only on request (tests/test_gpu_umma.py, MGN_TEST_PAIR=1) */
int mgn_selftest_umma_pair(const void* a, const void* b, float* d, int n, int ksteps, int a_tmem, void* stream);
/* Developer micro-benchmark (csrc/probe.cu, tools/probe.py): cycle counts of TMEM loads / stores by number of active
 * warps and of product issue-to-completion latencies on one SM; out_dev [64] int64. */
int mgn_probe_sm(long long* out_dev, int iters, void* stream);
/* Developer micro-benchmark of the SM <-> L2 / DRAM path with bulk copies (csrc/probe.cu: modes 0..6); tools/probe.py. */
int mgn_probe_mem(void* buf, const void* table, long long* out_dev, int mode, int chunks, int grid, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* MGN_B200_H */
