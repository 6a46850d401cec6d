/*
 * aether_b200 -- C ABI of the B200 (sm_100a) implementation of Aether's
 * state-to-state step.
 *
 * The reference is pure Python/PyTorch and has no FFI of its own; the entry
 * points below are what a binding for the hot path would bind, one per stage
 * the reference runs from Python.  Paths are relative to the reference repo
 * (mkofinas/aether):
 *
 *   aether_graph_prepare      replaces the per-batch edge handling of
 *                             experiments/lorentz/main.py:211-212 plus the index
 *                             work inside torch_scatter.scatter at
 *                             nn/state2state/locs/locs.py:236-238 (the reference
 *                             scatters unsorted; this library sorts by receiver
 *                             once and reuses the CSR in every layer).
 *   aether_fc_edges           experiments/lorentz/dataset4newton.py:54-61,84-94
 *   aether_knn_edges          experiments/ind/single_ind_data.py:186-217
 *   aether_caller_prep        experiments/lorentz/main.py:243-246
 *   aether_step_forward       Aether.forward, nn/state2state/aether.py:169-186
 *                             (FieldNetwork :127-134, AetherLocalizer :33-100,
 *                             GNN locs.py:183-194, GNNLayer locs.py:227-243,
 *                             Globalizer nn/utils/local_to_global.py:12-13)
 *   aether_step_backward      loss.backward() through that forward,
 *                             experiments/lorentz/main.py:291
 *   aether_rollout_step       one harness-level rollout step:
 *                             caller prep + forward + v' = x' - x
 *   aether_small_rollout      T such steps of a batch of small scenes in ONE
 *                             launch (the eval loops of experiments/lorentz/
 *                             main.py:243-247 and, with per-step kNN edges,
 *                             experiments/ind/single_ind_data.py:186-217)
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless the name ends in _host;
 *   - the caller owns all memory, including the workspace; sizes come from
 *     aether_workspace_bytes / aether_graph_scratch_bytes;
 *   - all work is enqueued on the stream passed in (a cudaStream_t cast to
 *     void*); no call synchronises or reads device data on the host, so every
 *     call can be captured in a CUDA graph.  The only process-wide state is
 *     diagnostic: an atomic launch counter (aether_launch_count) and the
 *     mutex-guarded event list of aether_profile_enable / _collect;
 *   - return value: 0 on success, otherwise a cudaError_t value or one of the
 *     negative AETHER_E_* codes; nothing throws or aborts.
 *   - floating point is fp32 at this boundary; indices are int64 on input (as
 *     the reference passes them) and int32 inside the prepared graph.
 */
#ifndef AETHER_B200_H_
#define AETHER_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define AETHER_B200_ABI_VERSION 1

/* error codes (negative; positive values are cudaError_t) */
#define AETHER_E_BADARG   (-1)
#define AETHER_E_WORKSPACE (-2)
#define AETHER_E_UNSUPPORTED (-3)
#define AETHER_E_TOO_LARGE (-4)   /* beyond the fixed maximum of an entry point that works on static scratch memory */

/* flags for aether_step_forward / aether_workspace_bytes */
#define AETHER_F_SAVE_FOR_BACKWARD 1u   /* keep activations for aether_step_backward  */
#define AETHER_F_BF16_EDGE_MLP     2u   /* edge-MLP GEMMs on tcgen05 in bf16 (fp32 accumulate) */
#define AETHER_F_EXTERNAL_FIELD    4u   /* the field at the particles is an INPUT (DynamicFieldAether,
                                         * nn/state2state/dynamic_field_aether.py:84-88): params[40] is a device pointer
                                         * to field[Nn, D] instead of field_net.net.0.weight, params[41..46] and
                                         * `charges` are not read; aether_step_backward writes d loss / d field to
                                         * param_grads[40] ([Nn, D]) and leaves param_grads[41..46] untouched */
/* (d/dx and d/dvel are returned by aether_step_backward whenever gx_out / gvel_out are non-null) */

/* Number of parameter tensors of Aether, in state_dict order
 * (nn/state2state/aether.py:143-159, locs.py:160-181):
 *   0..5    gnn.out_mlp.{0,3,6}.{weight,bias}
 *   6..15   gnn.layer_1.{message_fn.0, message_fn.2, res, update_fn.0, update_fn.2}.{weight,bias}
 *   16..23  gnn.layer_2.{message_fn.0, message_fn.2, update_fn.0, update_fn.2}.{weight,bias}
 *   24..31  gnn.layer_3.(same)        32..39  gnn.layer_4.(same)
 *   40..45  field_net.net.{0,2,4}.{weight,bias}
 *   46      field_net.class_embedding.weight
 */
#define AETHER_NUM_PARAMS 47

typedef struct AetherDims {
  int32_t num_dims;     /* D: 2 or 3                                   */
  int32_t hidden;       /* H: multiple of 4                            */
  int64_t n_nodes;      /* Nn: nodes flattened over the batch          */
  int64_t n_edges;      /* E                                           */
} AetherDims;

/* Receiver-sorted view of an edge list, all arrays caller-allocated:
 *   perm[E]      sorted position -> original edge id (stable by receiver)
 *   rowptr[Nn+1] segment offsets of receivers in the sorted order
 *   send[E], recv[E]   endpoints in sorted order
 *   sperm[E]     positions of the sorted list grouped by sender (stable)
 *   colptr[Nn+1] segment offsets of senders in sperm
 */
typedef struct AetherGraph {
  int32_t* perm;
  int32_t* rowptr;
  int32_t* send;
  int32_t* recv;
  int32_t* sperm;
  int32_t* colptr;
} AetherGraph;

int aether_abi_version(void);
const char* aether_error_string(int code);

/* ---- integer path -------------------------------------------------------- */
size_t aether_graph_scratch_bytes(int64_t n_nodes, int64_t n_edges);
int aether_graph_prepare(const int64_t* send, const int64_t* recv, int64_t n_edges, int64_t n_nodes,
                         AetherGraph* graph, int32_t* status /* device, 1 word: !=0 on out-of-range index */,
                         void* scratch, size_t scratch_bytes, void* stream);
/* aether_graph_prepare for callers that cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops,
 * harnesses that forbid it).  NO COUNTERPART IN THE REFERENCE.  Same
 * inputs, outputs and results as aether_graph_prepare, but the scratch is a static __device__ array inside the library
 * (2.5 MB, held by every GPU that loads the library until the process exits), so neither the caller nor the library
 * allocates anything: no cudaMalloc / cudaFree / allocating library routine is called.
 *   Limits: n_nodes <= AETHER_GRAPH_STATIC_MAX_NODES and n_edges <= AETHER_GRAPH_STATIC_MAX_EDGES, otherwise
 *   AETHER_E_TOO_LARGE is returned and nothing is launched (use aether_graph_prepare with your own scratch).
 *   Every call on a device shares the one array, which keeps the previous call's contents (the call clears what it
 *   needs at zero itself): calls that overlap on the same GPU -- different streams, or two host threads -- RACE; run
 *   them on one stream or order them with events.  The first call on a device looks up the array's address
 *   (cudaGetSymbolAddress, that device current); make that first call outside stream capture. */
#define AETHER_GRAPH_STATIC_MAX_NODES 65536
#define AETHER_GRAPH_STATIC_MAX_EDGES 524288
int aether_graph_prepare_static(const int64_t* send, const int64_t* recv, int64_t n_edges, int64_t n_nodes,
                                AetherGraph* graph, int32_t* status /* device, 1 word */, void* stream);
int aether_fc_edges(int64_t nodes_per_graph, int64_t n_graphs, int64_t* send, int64_t* recv, void* stream);
/* Ragged batch of scenes; scene s owns nodes [node_off[s], node_off[s+1]).  Writes
 * edge_off[s] + i*kk + j for query agent i of scene s (kk = min(k, n_s - 1)).  edge_off is
 * computed by the caller (it depends only on the scene sizes). */
int aether_knn_edges(const float* pos, int32_t pos_stride, const int32_t* node_off, const int64_t* edge_off,
                     int32_t n_scenes, int32_t k, int64_t* send, int64_t* recv, void* stream);
int aether_caller_prep(const float* x, const float* vel, const float* charges, int32_t num_dims,
                       const int64_t* send, const int64_t* recv, int64_t n_edges, int64_t n_nodes,
                       float* h /* [Nn] or NULL */, float* edge_attr_orig /* [E,2] */, void* stream);

/* Radius graph capped at the k nearest (k <= 32), for single scenes too large for aether_knn_edges (BASELINE config 5;
 * the reference only has the O(N^2) cdist + topk of experiments/ind/single_ind_data.py:186-217).  For every sender i the
 * receivers are the particles j != i with |x_i - x_j|^2 <= radius^2, nearest first, ties toward the lower index;
 * sender-major like get_knn_graph_info.  Uniform cell list: cell edge cell_size >= radius (a hair larger, so that
 * rounding never puts a neighbour two cells away), box_lo = lower corner, grid_dim[a] = cells along axis a (host arrays).
 *   _neighbors: nbr[n,k] (-1 padded), off[n+1] = exclusive scan of the degrees (off[n] = E).  The caller reads off[n]
 *               on the host (the one synchronisation of graph construction), allocates send/recv[E], then
 *   _fill:      writes the edge list. */
size_t aether_radius_graph_scratch_bytes(int64_t n, int32_t num_dims, const int32_t* grid_dim);
int aether_radius_graph_neighbors(const float* pos, int32_t pos_stride, int32_t num_dims, int64_t n, float radius,
                                  float cell_size, int32_t k, const float* box_lo /* host [num_dims] */,
                                  const int32_t* grid_dim /* host [num_dims] */, int32_t* nbr, int32_t* off,
                                  void* scratch, size_t scratch_bytes, void* stream);
int aether_radius_graph_fill(const int32_t* nbr, const int32_t* off, int64_t n, int32_t k, int64_t* send, int64_t* recv,
                             void* stream);

/* ---- the step ------------------------------------------------------------ */
size_t aether_workspace_bytes(const AetherDims* dims, uint32_t flags);
size_t aether_backward_scratch_bytes(const AetherDims* dims, uint32_t flags);

int aether_step_forward(const AetherDims* dims, const float* const* params /* [47] device ptrs, host array */,
                        const AetherGraph* graph, const float* x, const float* vel, const float* charges,
                        const float* edge_attr_orig /* [E,2], ORIGINAL edge order */,
                        const float* dropout_mask /* [2,Nn,H] scale-or-zero, or NULL */,
                        float* out /* [Nn,D] */, void* workspace, size_t workspace_bytes, uint32_t flags,
                        void* stream);

int aether_step_backward(const AetherDims* dims, const float* const* params, const AetherGraph* graph,
                         const float* x, const float* vel, const float* charges, const float* edge_attr_orig,
                         const float* dropout_mask, const float* grad_out /* [Nn,D] */,
                         float* const* param_grads /* [47] device ptrs, overwritten */,
                         float* grad_x /* [Nn,D] or NULL */, float* grad_vel /* [Nn,D] or NULL */,
                         void* workspace /* the one forward filled */, size_t workspace_bytes,
                         void* scratch, size_t scratch_bytes, uint32_t flags, void* stream);

/* Receiver-partitioned inference step for ONE large scene.  This rank owns the node rows
 * [own_begin, own_end) and every edge whose receiver it owns: `graph` holds only those edges, receiver-sorted, with
 * GLOBAL node ids in send/recv, rowptr[n_nodes+1] indexing the local edge array (empty segments outside the owned
 * rows) and perm mapping local sorted position -> row of edge_attr_orig; sperm/colptr are not used.
 * dims->n_nodes = nodes of the whole scene, dims->n_edges = local edges.  x / vel / charges hold ALL nodes.
 * The step is cut where node state must cross ranks:
 *   phase 0: field + frames (all nodes), features + layer 1 (local edges, owned rows), sender terms of layer 2
 *   phase 1, 2: layer phase+1, sender terms of the next layer
 *   phase 3: layer 4, out MLP, rotation back, Euler update -> out rows [own_begin, own_end)
 * Phases 0-2 end by writing rows [own_begin, own_end) of xchg_out ([>= n_nodes, hidden], fp32, or bf16 with
 * AETHER_F_BF16_EDGE_MLP): the sender term P_s = x W_s^T of the node-factorised first Linear.  The caller
 * all-gathers it (NCCL, torch.distributed) and passes it as xchg_in of the next phase, whose edge GEMM gathers
 * rows of it by sender.  xchg_in and xchg_out may be the same buffer when every rank has its own (the phase reads
 * before it writes); partitions that share one address space must alternate two buffers.
 * The workspace (aether_workspace_bytes of the same dims, without SAVE_FOR_BACKWARD) must persist across the phases. */
int aether_part_forward_phase(const AetherDims* dims, const float* const* params, const AetherGraph* graph,
                              const float* x, const float* vel, const float* charges, const float* edge_attr_orig,
                              int64_t own_begin, int64_t own_end, int32_t phase, const void* xchg_in, void* xchg_out,
                              float* out,
                              void* workspace, size_t workspace_bytes, uint32_t flags, void* stream);

/* The exchange between two phases without NCCL (csrc/peer.cu; SURVEY.md 8e "producer writes its shard directly into
 * peers' buffers over NVLink"; the reference is single-device and has no collective to replace).  Every rank holds the
 * same symmetric allocation; peer_base[r] (HOST array of `world` device pointers) is rank r's copy as mapped into this
 * process.  The buffer at buf_offset is `world` blocks of block_bytes; this rank has written block `rank` of its own
 * copy.  One launch stores that block into every peer's copy, raises flags[rank] = epoch in every peer's flag row
 * (world uint32 at flags_offset of the allocation, zero before the first call) and waits until its own row shows the
 * epoch from every peer -- the usual in-place all-gather, done by the producers.  state: 4 uint32 of LOCAL device
 * memory, zero before the first call ([0] epoch, [1] internal, [2] != 0 after a wait that gave up because a peer never
 * arrived).  All ranks must issue the same sequence of calls.  Offsets and sizes are multiples of 4 (16 for vector
 * stores); world <= 16. */
int aether_peer_allgather(void* const* peer_base, int32_t world, int32_t rank, size_t buf_offset, size_t block_bytes,
                          size_t flags_offset, uint32_t* state, void* stream);

/* One harness-level rollout step: edge_attr = [q_i q_j, |x_i - x_j|] over the ORIGINAL
 * edge list (send64/recv64, experiments/lorentz/main.py:243-246), x_next = step(x, vel),
 * vel_next = x_next - x.  x_next/vel_next must not alias x/vel. */
int aether_rollout_step(const AetherDims* dims, const float* const* params, const AetherGraph* graph,
                        const int64_t* send64, const int64_t* recv64, const float* x, const float* vel,
                        const float* charges, float* x_next, float* vel_next, float* edge_attr_scratch /* [E,2] */,
                        void* workspace, size_t workspace_bytes, uint32_t flags, void* stream);

/* T rollout steps of a batch of small scenes in one launch (csrc/small.cu): one CTA keeps a whole scene -- edge state,
 * node state, frames, edge list -- in shared memory for all steps; per step it does what aether_rollout_step does
 * (inference only).  flags = 0: fp32 arithmetic on the CUDA cores (the fp32 tolerance); AETHER_F_BF16_EDGE_MLP: the
 * edge MLP on the tensor cores (mma.sync, bf16 operands, fp32 accumulate; edge state kept as bf16 in shared memory:
 * the bf16 tolerance, twice the edges per scene), everything else still fp32.  A small pack
 * launch in front writes the step's 32 weight tiles (64 x 64, transposed) into `scratch`
 * (aether_small_rollout_scratch_bytes, 128-byte aligned); the rollout kernel streams them from there (L2) by bulk copies.
 *   node_off [n_scenes+1] (int32), edge_off [n_scenes+1] (int64): first particle / first edge of every scene;
 *   order [n_scenes] (int32, optional): the scene CTA i works on -- largest scenes first balances the SMs;
 *   send/recv: int64 edge list with global particle ids whose edges are grouped by scene (edge_off) and stay inside
 *   their scene, or both NULL: the knn_k nearest neighbours of every agent are rebuilt from the positions every step
 *   (edge_off then has to be the kNN plan's: n_s * min(knn_k, n_s - 1) edges per scene);
 *   max_nodes / max_edges: upper bounds over the scenes (host-known; they size the shared memory);
 *   traj [steps, n_nodes, D]: positions after every step; send_out/recv_out (optional, kNN): the last step's edge list.
 * aether_small_rollout_supported says whether a shape fits (hidden 64, <= 64 particles per scene, edge state plus a
 * three-tile weight ring within one SM's shared memory); unsupported shapes -- and parameter tensors that are not
 * 16-byte aligned -- return AETHER_E_UNSUPPORTED and the caller uses aether_rollout_step. */
size_t aether_small_rollout_scratch_bytes(void);
int aether_small_rollout_supported(int32_t num_dims, int32_t hidden, int32_t max_nodes, int32_t max_edges, uint32_t flags);
int aether_small_rollout(const AetherDims* dims, const float* const* params, int32_t n_scenes, const int32_t* node_off,
                         const int64_t* edge_off, const int32_t* order, const int64_t* send, const int64_t* recv,
                         int32_t knn_k, int32_t max_nodes, int32_t max_edges, const float* x0, const float* v0,
                         const float* charges, int32_t steps, float* traj, int64_t* send_out, int64_t* recv_out,
                         void* scratch, size_t scratch_bytes, uint32_t flags, void* stream);

/* Bench support: kernels launched by this library so far; optional per-category CUDA-event timing
 * (categories: 0 edge-MLP forward, 1 edge backward data grads, 2 edge weight grads, 3 segment
 * reductions of the backward pass, 4 edge features (gather + frame transform), 5 node-level work, 6 receiver mean of
 * the forward pass).  collect() synchronises the device. */
uint64_t aether_launch_count(void);
int aether_profile_enable(int on);
int aether_profile_collect(double* ms_by_cat, int64_t* spans_by_cat, int32_t n_cat);

/* Debug/test hook: copy a named stage out of a forward workspace (fp32, device to device).
 * Stages: "field" [Nn,D], "rel_feat" [Nn,3D], "R" [Nn,D,D], "edge_attr" [E,Fe] (sorted order),
 * "e1".."e4" [E,H] (sorted order), "x1".."x4" [Nn,H], "pred" [Nn,D].  Returns the number of
 * floats written, or a negative error. */
int64_t aether_debug_stage(const AetherDims* dims, const void* workspace, uint32_t flags, const char* stage,
                           float* dst, int64_t dst_floats, void* stream);

/* ---- latent field network of DynamicFieldAether (nn/state2state/dynamic_field_aether.py:11-47; csrc/dynfield.cu) -------
 * field[Nn, D] = FilmedNetwork([x, v, emb(q)], z[graph]) with z = attention pooling of the graph's nodes
 * (graph_pool.py:6-29, film.py:5-60).  Graphs are blocks of num_nodes consecutive nodes (Nn % num_nodes == 0).
 * params / param_grads: AETHER_DYNFIELD_NUM_PARAMS fp32 tensors in the state_dict order of `field_net`:
 *   0..3   summary_net.summary_net.gate_nn.{0,2}.{weight,bias}     4..7   summary_net.summary_net.nn.{0,2}.{weight,bias}
 *   8..13  wrapper.linear_{1,2,3}.{weight,bias}
 *   14..19 wrapper.film_1.modulator.{0,2,4}.{weight,bias}          20..25 wrapper.film_2.modulator.(same)
 *   26     class_embedding.weight
 * The step itself then runs with AETHER_F_EXTERNAL_FIELD.  aether_dynfield_backward takes d loss / d field (what
 * aether_step_backward left in param_grads[40]), writes every parameter gradient and ADDS the field's dependence on the
 * inputs to grad_x / grad_vel (either may be NULL). */
#define AETHER_DYNFIELD_NUM_PARAMS 27
size_t aether_dynfield_workspace_bytes(int32_t num_dims, int64_t n_nodes, int32_t num_nodes);
size_t aether_dynfield_scratch_bytes(int32_t num_dims, int64_t n_nodes, int32_t num_nodes);
int aether_dynfield_forward(int32_t num_dims, int64_t n_nodes, int32_t num_nodes, const float* const* params,
                            const float* x, const float* vel, const float* charges, float* field, void* workspace,
                            size_t workspace_bytes, void* stream);
int aether_dynfield_backward(int32_t num_dims, int64_t n_nodes, int32_t num_nodes, const float* const* params,
                             const float* charges, const float* grad_field, float* const* param_grads, float* grad_x,
                             float* grad_vel, const void* workspace, size_t workspace_bytes, void* scratch,
                             size_t scratch_bytes, void* stream);

/* ---- seq2seq decoder path (aether_b200/csrc/seqdec.cu) ---------------------------------------------------------------
 * The paper's rollout model for the electrostatic / gravitational experiments: one step of
 * RecurrentDecoder.forward (nn/seq2seq/aether.py:589-654; per-edge-type message MLPs weighted by the edge-type
 * weights, receiver mean, AugmentedLocalizer nn/utils/augmented_global_to_local.py:51-68 with its virtual origin node,
 * GRU-style gates, out MLP, Globalizer, inputs + prediction) and the field predictor of the enclosing model
 * (Aether.predict_field, nn/seq2seq/aether.py:83-87: Fourier features nn/nn/fourier_feature_mapper.py:21-23 + field_net
 * :72-78).  Inference only (dropout inactive).  The graph is the reference's fixed one: num_vars particles per
 * trajectory, every ordered pair (sender-major, torch.where(~eye), :548-549).  The decoder takes the
 * edge-type weights as an input; aether_seq_encoder_step below is the prior branch of the encoder that produces their logits.
 *
 * Layouts: states[B, N, 2D] = [pos | vel]; hidden[B, N, H]; edge_types[B, N(N-1), n_types] (leading dimension
 * n_types; soft weights or one-hot); field[B, N, D].  hidden_out may alias hidden_in.
 * Every pointer of AetherSeqParams is nn.Linear layout ([out, in] row-major weight, [out] bias), named after the
 * reference's state_dict keys; per-type arrays are indexed by edge type (entries below `skip_first` are not read).
 * The field_* / rff_B members are only read by aether_seq_predict_field / aether_seq_rollout_step. */
#define AETHER_SEQ_MAX_TYPES 4
/* every Linear on the tensor cores (mma.sync, bf16 operands rounded from the fp32 tensors on the fly, fp32 accumulate;
 * biases, activations, sums and all inputs / outputs stay fp32): the bf16 tolerance (2e-2) instead of 1e-5 */
#define AETHER_SEQ_F_BF16 1u
/* every Linear on the tensor cores with fp32 accuracy: each fp32 operand split into three bf16 (hi + mid + lo, 24 mantissa
 * bits), six mma.sync products per operand pair, fp32 accumulate -- the fp32 tolerance (1e-5), not bit-identical to the
 * CUDA-core summation order */
#define AETHER_SEQ_F_SPLIT 2u
typedef struct AetherSeqDims {
  int32_t num_dims;      /* D: 2 or 3 (use_3d)                              */
  int32_t num_vars;      /* N particles per trajectory, 2..64               */
  int32_t hidden;        /* decoder_hidden, multiple of 4                   */
  int32_t n_types;       /* num_edge_types, 1..AETHER_SEQ_MAX_TYPES         */
  int32_t skip_first;    /* 1: edge type 0 sends no message                 */
  int32_t field_hidden;  /* encoder_hidden (width of field_net and of the encoder), multiple of 4; 0 when unused */
  int32_t rnn_hidden;    /* encoder_rnn_hidden (LSTM width); 0 when aether_seq_encoder_step is not used */
  int32_t prior_hidden;  /* prior_hidden_size (0 with a one-layer prior head)                            */
  uint32_t flags;        /* AETHER_SEQ_F_*                                                               */
  int64_t batch;         /* B trajectories                                  */
} AetherSeqDims;
typedef struct AetherSeqParams {
  const float* msg_fc1_w[AETHER_SEQ_MAX_TYPES];          /* [H, 2H]   (columns: receiver state | sender state) */
  const float* msg_fc1_b[AETHER_SEQ_MAX_TYPES];
  const float* msg_fc2_w[AETHER_SEQ_MAX_TYPES];          /* [H, H] */
  const float* msg_fc2_b[AETHER_SEQ_MAX_TYPES];
  const float* hidden_r_w; const float* hidden_i_w; const float* hidden_h_w;             /* [H, H], no bias */
  const float* present_r_w; const float* present_r_b;    /* [H, H] */
  const float* present_i_w; const float* present_i_b;
  const float* present_n_w; const float* present_n_b;
  const float* out_w[3]; const float* out_b[3];          /* out_mlp.{0,3,6}: [H,H], [H,H], [2D,H] */
  const float* present_msg_fc1_w[AETHER_SEQ_MAX_TYPES];  /* [H, 2F+3D], F = 4D + D(D-1)/2 */
  const float* present_msg_fc1_b[AETHER_SEQ_MAX_TYPES];
  const float* present_msg_fc2_w[AETHER_SEQ_MAX_TYPES];  /* [H, H] */
  const float* present_msg_fc2_b[AETHER_SEQ_MAX_TYPES];
  const float* input_r_w; const float* input_r_b;        /* [H, 3D+F] */
  const float* input_i_w; const float* input_i_b;
  const float* input_n_w; const float* input_n_b;
  const float* rff_B;                                    /* coordinate_embedding.B: [D, field_hidden/2] */
  const float* field_w[3]; const float* field_b[3];      /* field_net.{0,2,4}: [Hf,Hf], [Hf,Hf], [D,Hf] */
} AetherSeqParams;
size_t aether_seq_workspace_bytes(const AetherSeqDims* dims);
/* field[B,N,D] = field_net(rff(states[..., :D])) */
int aether_seq_predict_field(const AetherSeqDims* dims, const AetherSeqParams* params, const float* states, float* field,
                             void* workspace, size_t workspace_bytes, void* stream);
/* (states_out, hidden_out) = RecurrentDecoder.forward(states, hidden_in, edge_types, field) */
int aether_seq_decoder_step(const AetherSeqDims* dims, const AetherSeqParams* params, const float* states,
                            const float* hidden_in, const float* edge_types, const float* field, float* states_out,
                            float* hidden_out, void* workspace, size_t workspace_bytes, void* stream);
/* One step of the loops of Aether.predict_future (nn/seq2seq/aether.py:166-184): field = predict_field(states), then the
 * decoder step with it.  field_out (may be NULL) receives the predicted field. */
int aether_seq_rollout_step(const AetherSeqDims* dims, const AetherSeqParams* params, const float* states,
                            const float* hidden_in, const float* edge_types, float* states_out, float* hidden_out,
                            float* field_out, void* workspace, size_t workspace_bytes, void* stream);
/* Prior branch of the edge-type encoder, one time step (Encoder.single_step_forward, nn/seq2seq/aether.py:397-425; the
 * burn-in pass Encoder.forward :362-395 is the same step applied to every burn-in time step, eval mode): localizer ->
 * anisotropic edge filter (nn/nn/anisotropic_filter.py:12-41) -> edge2node + res1 -> mlp3 -> node2edge + skip -> mlp4
 * (RefNRIMLP, nn/utils/model_utils.py:15-54, BatchNorm in eval mode) -> one LSTM step -> prior_fc_out.
 *   h_in / c_in / h_out / c_out: [B*N(N-1), rnn_hidden] (out may alias in); logits: [B, N(N-1), n_types].
 * BatchNorm enters as its eval-mode affine map: *_bn_scale = weight / sqrt(running_var + eps), *_bn_shift = bias -
 * running_mean * scale ([encoder_hidden] each; NULL without batch norm).  pos_offset: first column of the edge features
 * that the filter generator reads (pos_representation 'cart': 0; 'polar': 2 in 2D, 3 in 3D;
 * nn/utils/augmented_global_to_local.py:18-24), D + D(D-1)/2 columns. */
typedef struct AetherSeqEncParams {
  const float* filter_w0; const float* filter_b0;      /* edge_filter.edge_filter.0: [He, D+O]                      */
  const float* filter_w2; const float* filter_b2;      /* edge_filter.edge_filter.2: [(2F+3D)*He, He]               */
  const float* res1_w; const float* res1_b;            /* [He, 3D+F]                                                */
  const float* mlp3_w[2]; const float* mlp3_b[2];      /* mlp3.model.{0,3}: [He, He]                                */
  const float* mlp3_bn_scale; const float* mlp3_bn_shift;
  const float* mlp4_w[2]; const float* mlp4_b[2];      /* mlp4.model.{0,3}: [He, 3He], [He, He]                     */
  const float* mlp4_bn_scale; const float* mlp4_bn_shift;
  const float* lstm_w_ih; const float* lstm_w_hh;      /* forward_rnn.weight_{ih,hh}_l0: [4R, He], [4R, R]          */
  const float* lstm_b_ih; const float* lstm_b_hh;
  const float* prior_w[4]; const float* prior_b[4];    /* prior_fc_out Linears in order                             */
  int32_t n_prior_layers;                              /* prior_num_layers, 1..4                                    */
  int32_t pos_offset;
} AetherSeqEncParams;
int aether_seq_encoder_step(const AetherSeqDims* dims, const AetherSeqEncParams* params, const float* states,
                            const float* field, const float* h_in, const float* c_in, float* logits, float* h_out,
                            float* c_out, void* workspace, size_t workspace_bytes, void* stream);
/* Hard gumbel-softmax sample of the edge types (nn/utils/model_utils.py:57-121 with hard=True, as predict_future samples,
 * nn/seq2seq/aether.py:91-95): onehot[r, :] = one-hot of argmax_k(logits[r, k] - log(1e-10 - log(uniform[r, k] + 1e-10))),
 * first maximum on ties.  uniform: U(0,1) noise from the caller's generator, [rows, n_types] like logits. */
int aether_seq_sample_hard(const float* logits, const float* uniform, float* onehot, int64_t rows, int32_t n_types,
                           void* stream);
/* Test hook: the three AugmentedLocalizer outputs the step consumes, rel_feat[B,N,3D+F], Rinv[B,N,D,D],
 * edge_attr[B,N(N-1),2F+3D]. */
int aether_seq_localizer(const AetherSeqDims* dims, const float* states, const float* field, float* rel_feat, float* Rinv,
                         float* edge_attr, void* stream);

/* ---- optimizer step of the reference's caller (torch.optim.AdamW, experiments/lorentz/main.py:164,292) ----------
 * One fused update of n_tensors parameter tensors (host arrays of device pointers; numel[i] elements each, n_tensors <= 64):
 *   p <- p (1 - lr wd);  m <- b1 m + (1-b1) g;  v <- b2 v + (1-b2) g^2;  p <- p - lr/(1-b1^t) * m / (sqrt(v)/sqrt(1-b2^t) + eps)
 * step_dev: ONE float on the device holding the step count t; it is advanced by 1 before the update (a 1-thread launch),
 * so the call is capturable in a CUDA graph.  Same semantics and operation order as torch.optim.AdamW (not amsgrad,
 * not maximize). */
int aether_adamw_step(int32_t n_tensors, float* const* params, const float* const* grads, float* const* exp_avg,
                      float* const* exp_avg_sq, const int64_t* numel, float* step_dev, double lr, double beta1,
                      double beta2, double eps, double weight_decay, void* stream);

/* ---- test hooks for the tcgen05 building blocks (aether_b200/csrc/tc_gemm.cu, tc_wgrad.cu) -------
 * out[M,N] = epi(tf(A_bf16[M,K]) B^T), B[n][k] = transpose ? W[k*ldw+n] : W[n*ldw + (kmod ? k%kmod : k)],
 * zero outside [n_valid,k_valid]; epi adds bias[n]; with the bf16 output (exactly one of out_bf16 / out_f32
 * is given, leading dimension N) it also adds gat_bf16[idx_a[m]*ldg+n] + gat_bf16[idx_b[m]*ldg+goff_b+n] and
 * multiplies by silu'(dsilu_bf16[m,n]).
 * wpack_scratch: N*K*2 bytes. */
int aether_dbg_tc_gemm(const void* A_bf16, int64_t M, int32_t K, int64_t lda, const float* W, int64_t ldw,
                       int32_t N, int32_t n_valid, int32_t k_valid, int32_t transpose, int32_t kmod,
                       int32_t silu_on_load, const float* bias, const void* gat_bf16, int64_t ldg,
                       const int32_t* idx_a, const int32_t* idx_b, int32_t goff_b, const void* dsilu_bf16,
                       void* out_bf16, float* out_f32, void* wpack_scratch, void* stream);
/* dW[MG, n_valid] = sum_e G[e,:]^T tf(ACT[e,:]) (fold32: columns j and 32+j summed);
 * db[MG] = column sums of G when db is not NULL. */
size_t aether_dbg_tc_wgrad_partial_floats(int32_t MG, int32_t NA);
int aether_dbg_tc_wgrad(const void* G_bf16, int32_t MG, const void* ACT_bf16, int32_t NA, int32_t silu_on_load,
                        int64_t E, float* dW, int32_t n_valid, int32_t fold32, float* db, float* partial,
                        size_t partial_floats, void* stream);
/* Chained edge MLP of one layer on a CTA pair (csrc/tc_chain.cu; GNNLayer.message_fn, locs.py:206-212,235):
 * out1[M,H] = tf(A[M,K1]) W1^T + bias1 (+ gat_bf16[idx_a[m]*ldg+n] + gat_bf16[idx_b[m]*ldg+goff_b+n]),
 * out2[M,H] = silu(out1) W2^T + bias2, both bf16; out1_bf16 may be NULL (intermediate not stored).
 * W1 is [H, ldw1] fp32 (columns < k1_valid used, column k read at k % kmod1 when kmod1 != 0), W2 is [H, H].
 * wpack_scratch: (K1 + H) * H * 2 bytes. */
int aether_dbg_tc_chain(const void* A_bf16, int64_t M, int32_t K1, int64_t lda, const float* W1, int64_t ldw1,
                        int32_t k1_valid, int32_t kmod1, const float* W2, int32_t H, int32_t silu_on_load,
                        const float* bias1, const float* bias2, const void* gat_bf16, int64_t ldg,
                        const int32_t* idx_a, const int32_t* idx_b, int32_t goff_b, void* out1_bf16,
                        void* out2_bf16, void* wpack_scratch, void* stream);
/* The same with the sender term fed by TMA boxes (box_gather != 0; gat_rows = rows of the idx_a plane of gat_bf16): tiles
 * whose idx_a values span fewer than 128 rows read P_s from one 128-row box per 64 columns, the others load directly. */
int aether_dbg_tc_chain_ex(const void* A_bf16, int64_t M, int32_t K1, int64_t lda, const float* W1, int64_t ldw1,
                           int32_t k1_valid, int32_t kmod1, const float* W2, int32_t H, int32_t silu_on_load,
                           const float* bias1, const float* bias2, const void* gat_bf16, int64_t ldg,
                           const int32_t* idx_a, const int32_t* idx_b, int32_t goff_b, void* out1_bf16,
                           void* out2_bf16, void* wpack_scratch, int64_t gat_rows, int32_t box_gather, void* stream);
/* Chained data-gradient GEMMs of the edge-MLP backward on a CTA pair (csrc/tc_chain_bwd.cu; autograd of locs.py:232-235):
 * out1[M,H] = (A[M,H] W1 (+ gat_bf16[idx_b[m]*ldg + n])) * silu'(z1[m,n]),  out2[M,H] = (out1 W2) * silu'(z2[m,n]);
 * everything bf16 except W1, W2 ([H,H] fp32 row-major).  wpack_scratch: 2 * H * H * 2 bytes. */
int aether_dbg_tc_chain_bwd(const void* A_bf16, int64_t M, int32_t H, const float* W1, const float* W2,
                            const void* gat_bf16, int64_t ldg, const int32_t* idx_b, const void* z1_bf16,
                            const void* z2_bf16, void* out1_bf16, void* out2_bf16, void* wpack_scratch, void* stream);
/* Probe of the TMA row gather (tile::gather4): rows idx128[0..127] of a bf16 table, columns [col0, col0+64), into one
 * 128-byte-swizzled 16 KB operand slot; raw_out receives the slot's 16384 bytes. */
int aether_dbg_gather4_probe(const void* table_bf16, int64_t n_rows, int32_t n_cols, int64_t ld, int32_t box_rows,
                             const int32_t* idx128, int32_t col0, void* raw_out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* AETHER_B200_H_ */
