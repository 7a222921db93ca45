/* gnn_manip_b200 -- C ABI of the B200-native gnn-manip rollout hot path.
 *
 * Drop-in boundary for the Encode-Process-Decode GraphNet rollout of dblanm/gnn-manip.  The reference has no
 * FFI of its own (it is pure Python); each entry point below names the reference function it replaces
 * (paths relative to the reference checkout) and is what a ctypes / torch-extension binding on the
 * reference side calls.
 *
 * Conventions
 *   - every pointer is DEVICE memory unless the name ends in _host; tensors are dense, row-major, fp32 / int32
 *   - the library never allocates device memory: scratch is a caller-provided workspace whose size comes
 *     from the matching *_workspace_bytes() query
 *   - all work is enqueued on `stream` (a cudaStream_t passed as void*); nothing synchronises the host
 *   - return value: 0 = ok, otherwise a gm_status code; gm_last_error() gives the message (thread-local)
 *   - no exceptions cross the ABI
 *   - graph sizes that are only known on the device (edge counts) stay on the device: kernels are launched
 *     for the static upper bound max_edges = n_nodes * max_neighbours and read the live count themselves,
 *     which is what makes a whole rollout step capturable in a CUDA graph
 */
#ifndef GNN_MANIP_B200_H_
#define GNN_MANIP_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum gm_status {
  GM_STATUS_OK = 0,
  GM_STATUS_INVALID = 1,
  GM_STATUS_WORKSPACE = 2,
  GM_STATUS_CUDA = 3,
  GM_STATUS_UNSUPPORTED = 4
};

/* which kernels run the MLPs */
enum gm_impl {
  GM_IMPL_AUTO = 0,   /* tcgen05 path when the model is hidden=128 / num_layers=2 / LayerNorm, else fp32 CUDA cores */
  GM_IMPL_SIMT = 1,   /* generic fp32 CUDA-core kernels (any widths) */
  GM_IMPL_TC = 2      /* tcgen05 / TMEM kernels (error if the model shape is not covered) */
};

int gm_version(void);
const char* gm_last_error(void);
/* kernels of this library launched so far by the calling thread (launch-site counter; a replayed CUDA graph
 * counts gm_rollout_kernel_launches_per_step() per step instead) */
long long gm_launch_count(void);

/* Dataset constants (schema: gnn_manip/utils/coffee_dataset.py:18-43, read_metadata). Host struct, by pointer. */
typedef struct gm_phys {
  int32_t cartesian_idx[3];
  int32_t material_idx;
  int32_t control_idx[3];     /* ignored when use_control == 0 */
  int32_t use_control;        /* GraphBoundedMultimaterialControl vs GraphBoundedMultimaterial */
  float velocity_mean[3], velocity_std[3];
  float acceleration_mean[3], acceleration_std[3];
  float lower_bounds[3], upper_bounds[3];
  double conn_r;
} gm_phys;

/* ---------------------------------------------------------------------------------------------------------
 * A1  radius graph -- replaces get_connectivity (gnn_manip/utils/utils.py:64-93; sklearn KDTree.query_radius,
 *     sort_results=True, first max_neighbours kept per query node, self-loop included).
 *
 * pos: first coordinate of node 0; node i is at pos + i*pos_stride (x,y,z contiguous), so the last frame of an
 *      observation tensor obs[k,N,D] can be passed in place (pos = obs_last + cartesian_idx[0], stride = D).
 * n_graphs independent graphs of n_per_graph nodes each are built in one call; node ids are global
 *      (graph g owns [g*n_per_graph, (g+1)*n_per_graph)), the offset rule of collate_utils.py:75-76.
 * Outputs (int32):
 *   nbr [n, K] / cnt [n]   neighbour table; flattened row-major over valid entries it IS the reference edge
 *                          list: senders = repeat(i, cnt[i]), receivers = nbr[i, :cnt[i]]
 *   edge_ptr [n+1]         exclusive scan of cnt (reference edge id of (i, s) = edge_ptr[i] + s)
 *   recv_ptr [n+1], esend / erecv / eref [n*K]   the same edges sorted by (receiver, sender): CSR offsets,
 *                          sender, receiver and reference edge id of every sorted edge
 *   n_edges [1]            number of edges
 * --------------------------------------------------------------------------------------------------------- */
size_t gm_radius_graph_workspace_bytes(int32_t n_per_graph, int32_t n_graphs, int32_t max_neighbours);
int gm_radius_graph(const float* pos, int32_t pos_stride, int32_t n_per_graph, int32_t n_graphs, double conn_r,
                    int32_t max_neighbours, int32_t* nbr, int32_t* cnt, int32_t* edge_ptr, int32_t* recv_ptr,
                    int32_t* esend, int32_t* erecv, int32_t* eref, int32_t* n_edges, void* workspace,
                    size_t workspace_bytes, void* stream);

/* Expand (nbr, cnt, edge_ptr) into the reference's int64 edge_index [2, n_edges_host] (senders row 0,
 * receivers row 1) -- the compatibility surface of process_collate (collate_utils.py:68-87). */
int gm_edge_index_i64(const int32_t* nbr, const int32_t* cnt, const int32_t* edge_ptr, int32_t n_nodes,
                      int32_t max_neighbours, int64_t n_edges_host, int64_t* edge_index, void* stream);

/* Receiver CSR of an explicit edge list (any graph): stable sort of edges by receiver.
 * senders/receivers int32 [n_edges]; outputs as in gm_radius_graph. */
size_t gm_recv_csr_workspace_bytes(int32_t n_nodes, int64_t n_edges);
int gm_recv_csr(const int32_t* senders, const int32_t* receivers, int64_t n_edges, int32_t n_nodes,
                int32_t* recv_ptr, int32_t* esend, int32_t* erecv, int32_t* eref, void* workspace,
                size_t workspace_bytes, void* stream);

/* ---------------------------------------------------------------------------------------------------------
 * A2  features -- replaces GraphBoundedMultimaterial(Control).compute_nodes (collate_utils.py:195-232) with
 *     get_nodes_vel (utils.py:27-40), and get_edges_displacement (utils.py:43-61).
 * obs [k, n, D]; nodes [n, 3(k-1)+6+1(+3)]; edge_attr [n_edges, 4] in the order of the given senders/receivers.
 * n_edges_dev may be NULL (then max_edges edges are processed).
 * --------------------------------------------------------------------------------------------------------- */
int32_t gm_node_dim(int32_t k, int32_t use_control);
int gm_node_features(const float* obs, int32_t k, int32_t n, int32_t D, const gm_phys* phys_host, float* nodes,
                     void* stream);
int gm_edge_features(const float* pos, int32_t pos_stride, const int32_t* senders, const int32_t* receivers,
                     const int32_t* n_edges_dev, int32_t max_edges, double conn_r, float* edge_attr, void* stream);

/* ---------------------------------------------------------------------------------------------------------
 * A3/A4  model -- replaces EncProcDecGNN (gnn_manip/models/epd_gnn.py:11-105) and the blocks it is made of
 *        (deps/torch-graphnet/torch_graphnet/graph_networks.py:19-137, utils/gn_utils.py:7-64).
 *
 * flat_params: every *unique* parameter of the state dict, fp32, concatenated in this order
 *   encoder.phi_edge, encoder.phi_node, processor.0.phi_edge, processor.0.phi_node, ..., decoder
 *   and inside each Sequential: Linear weights [out,in] then bias, layer by layer, then LayerNorm weight, bias.
 * blob: device memory that receives the packed weights (size from gm_model_blob_bytes); must outlive the model.
 * --------------------------------------------------------------------------------------------------------- */
typedef struct gm_model gm_model;
size_t gm_model_num_params(int32_t node_dim, int32_t edge_dim, int32_t out_dim, int32_t hidden, int32_t num_layers,
                           int32_t m_steps);
size_t gm_model_blob_bytes(int32_t node_dim, int32_t edge_dim, int32_t out_dim, int32_t hidden, int32_t num_layers,
                           int32_t m_steps);
int gm_model_create(gm_model** model, const float* flat_params, size_t n_params, int32_t node_dim, int32_t edge_dim,
                    int32_t out_dim, int32_t hidden, int32_t num_layers, int32_t m_steps, void* blob,
                    size_t blob_bytes, void* stream);
void gm_model_destroy(gm_model* model);
int gm_model_supports_tc(const gm_model* model);

/* EncProcDecGNN.forward(nodes, edge_attr, edge_index) -> pred [n_nodes, out_dim] (epd_gnn.py:86-98).
 * The graph is given in receiver-major form (gm_radius_graph / gm_recv_csr outputs).
 * edge_attr: explicit edge features [n_edges, edge_dim] indexed by REFERENCE edge id (eref maps sorted edges to
 *            them), or NULL to build the displacement features on the fly from pos (edge_dim must be 4). */
size_t gm_forward_workspace_bytes(const gm_model* model, int32_t n_nodes, int32_t max_edges, int32_t impl);
int gm_model_forward(const gm_model* model, int32_t impl, const float* nodes, const float* edge_attr,
                     const float* pos, int32_t pos_stride, double conn_r, int32_t n_nodes, int32_t max_edges,
                     const int32_t* n_edges_dev, const int32_t* recv_ptr, const int32_t* esend,
                     const int32_t* erecv, const int32_t* eref, float* pred, void* workspace,
                     size_t workspace_bytes, void* stream);

/* The three stages of EncProcDecGNN.forward as separate calls, for a caller that keeps the reference's structure
 * (encoder -> for in_module in processor: _process -> decoder; gnn_manip/models/epd_gnn.py:86-105).  The latents (and, on
 * the tensor path, the projections Pr / Ps of the next edge step) live in `workspace` between the calls: same size and
 * layout as gm_model_forward (gm_forward_workspace_bytes with the same impl), same graph arrays in every call.
 *   gm_encode   epd_gnn.py:88     latent_node_0, latent_edge_0 (edge features from edge_attr or, if NULL, from pos)
 *   gm_mp_step  epd_gnn.py:99-105 one processor block incl. both residuals; step = 0 .. m_steps - 1, in order
 *   gm_decode   epd_gnn.py:96     pred [n_nodes, out_dim]; fp32 CUDA-core MLP on both paths (gm_model_forward fuses the
 *                                 decoder into the last node kernel instead) */
int gm_encode(const gm_model* model, int32_t impl, const float* nodes, const float* edge_attr, const float* pos,
              int32_t pos_stride, double conn_r, int32_t n_nodes, int32_t max_edges, const int32_t* n_edges_dev,
              const int32_t* recv_ptr, const int32_t* esend, const int32_t* erecv, const int32_t* eref, void* workspace,
              size_t workspace_bytes, void* stream);
int gm_mp_step(const gm_model* model, int32_t impl, int32_t step, int32_t n_nodes, int32_t max_edges,
               const int32_t* n_edges_dev, const int32_t* recv_ptr, const int32_t* esend, const int32_t* erecv,
               void* workspace, size_t workspace_bytes, void* stream);
int gm_decode(const gm_model* model, int32_t impl, int32_t n_nodes, int32_t max_edges, float* pred, void* workspace,
              size_t workspace_bytes, void* stream);

/* Same forward on the tcgen05 path with a CUDA event recorded on `stream` before every kernel; synchronises the
 * stream and returns per-kernel-class device time: ms_host[4] / launches_host[4] for
 * {edge encoder, node encoder, mp_edge (edge MLP + segmented sum), mp_node (node MLP + projections / decoder)}.
 * Measurement aid for bench.py's roofline (no reference counterpart). */
int gm_model_forward_profile(const gm_model* model, const float* nodes, const float* edge_attr, const float* pos,
                             int32_t pos_stride, double conn_r, int32_t n_nodes, int32_t max_edges,
                             const int32_t* n_edges_dev, const int32_t* recv_ptr, const int32_t* esend,
                             const int32_t* erecv, const int32_t* eref, float* pred, void* workspace,
                             size_t workspace_bytes, void* stream, float* ms_host, int32_t* launches_host);

/* Block-level entry points (any widths; fp32 CUDA cores) behind EdgeModel / NodeModel / GraphIndependent:
 * out[m] = MLP(concat_s src[s][idx[s] ? idx[s][m] : m]) with ReLU between layers, optional LayerNorm(eps 1e-5).
 * weights_t[l]: TRANSPOSED Linear weight [in_l, out_l]; all arrays of pointers are HOST arrays of device pointers. */
int gm_mlp_forward(int32_t n_layers, const float* const* weights_t_host, const float* const* biases_host,
                   const int32_t* dims_host /* n_layers+1 */, const float* ln_weight, const float* ln_bias,
                   int32_t n_src, const float* const* src_host, const int32_t* const* idx_host,
                   const int32_t* widths_host, int64_t n_rows, float* out, void* stream);
/* scatter_sum (gn_utils.py:60-64) as a deterministic receiver-sorted segmented sum:
 * agg[n] = sum_{t in [recv_ptr[n], recv_ptr[n+1])} src[eref[t]] in ascending t. */
int gm_segment_sum(const float* src, int32_t width, const int32_t* recv_ptr, const int32_t* eref, int32_t n_nodes,
                   float* agg, void* stream);

/* ---------------------------------------------------------------------------------------------------------
 * A5  integrator + rollout loop -- replaces get_position_from_prediction (gnn_manip/utils/rollout_utils.py:145-158)
 *     and the loop body of CMAESolver.cma_objective (gnn_manip/utils/traj_utils.py:119-152) == compute_rollout
 *     (rollout_utils.py:38-65), including compute_particles_tmatrix (traj_utils.py:167-194).
 * --------------------------------------------------------------------------------------------------------- */
/* next_pos [n,3] = pos[-1] + ((pos[-1] - pos[-2]) + (pred*acc_std + acc_mean)); obs is not modified. */
int gm_next_position(const float* obs, int32_t k, int32_t n, int32_t D, const gm_phys* phys_host, const float* pred,
                     float* next_pos, void* stream);

typedef struct gm_rollout gm_rollout;
typedef struct gm_rollout_desc {
  const gm_model* model;
  int32_t impl;               /* gm_impl */
  gm_phys phys;
  int32_t k, D;               /* obs is [k, n_graphs*n_per_graph, D] */
  int32_t n_per_graph, n_graphs;
  int32_t max_neighbours;
  int32_t n_rigid;            /* rigid ("cup") particles per graph */
  const int32_t* rigid_idx;   /* [n_rigid] local node ids (device) */
  int32_t traj_steps;         /* rollout steps covered by the control trajectory */
  const float* traj;          /* optional [n_graphs, traj_steps, n_rigid, 3] rigid positions per step, or NULL ... */
  const double* rot;          /* ... then rot/ty [n_graphs, traj_steps] (radians / metres) drive compute_particles_tmatrix */
  const double* ty;
  const float* rigid_init;    /* [n_rigid, 3] rigid particle positions of the initial state (used with rot/ty) */
  float ty_init[3];
  float* obs;                 /* state, advanced in place */
  float* pred;                /* optional [n, out_dim]: last normalised acceleration */
  int32_t use_cuda_graph;     /* capture one step and replay it */
} gm_rollout_desc;

size_t gm_rollout_workspace_bytes(const gm_rollout_desc* desc_host);
int gm_rollout_create(gm_rollout** plan, const gm_rollout_desc* desc_host, void* workspace, size_t workspace_bytes,
                      void* stream);
/* run `steps` rollout steps starting at the plan's current step index (reset with gm_rollout_reset).  A tensor-core
 * plan owns one extra CUDA stream and two events (no device memory): the node encoder of each step is forked onto
 * it beside the graph build and joined back into `stream` before the first message-passing step, so all work is
 * still ordered by `stream` from the caller's point of view. */
int gm_rollout_run(gm_rollout* plan, int32_t steps, void* stream);
int gm_rollout_reset(gm_rollout* plan, int32_t step, void* stream);
/* Measurement aid: run ONE rollout step un-captured and strictly serially on `stream` with a CUDA event between
 * phases; synchronises the stream.  ms_host / launches_host [10]: apply_control, node_features, cell grid build,
 * neighbour_kernel, edge_ptr scan + receiver CSR sort, enc_edge, enc_node, mp_edge (all steps), mp_node (all steps),
 * integrate.  The step advances the plan's state like gm_rollout_run(plan, 1). */
int gm_rollout_step_profile(gm_rollout* plan, void* stream, float* ms_host, int32_t* launches_host);
/* device pointers into the plan's workspace holding the graph of the most recent step (for parity tests) */
int gm_rollout_graph(const gm_rollout* plan, const int32_t** nbr, const int32_t** cnt, const int32_t** edge_ptr,
                     const int32_t** n_edges, const float** nodes);
int64_t gm_rollout_kernel_launches_per_step(const gm_rollout* plan);
void gm_rollout_destroy(gm_rollout* plan);

/* ---- Row N2: CMA-ES trajectory parametrisations (gnn_manip/utils/traj_utils.py) ------------------------------
 * kind GM_TRAJ_INCREMENTS  TrajectoryCMAsolver.interpolate_trajectory (traj_utils.py:206-228): x holds per-step
 *                          increments, scaled, clipped to +-max_rot / +-max_ty and accumulated from (rx_init, 0);
 *                          n_points = horizon - 1 parameters per curve.
 * kind GM_TRAJ_PCHIP       InterpolatedCMAsolver.interpolate_trajectory (traj_utils.py:393-416): x holds the knots
 *                          after the first (one every knot_stride steps, n_points * knot_stride == horizon), scaled
 *                          and offset from (rx_init, 0), interpolated like scipy.interpolate.pchip_interpolate.
 * x [n_cand, 2 * n_points] (rotation parameters first, then translation, as the reference's search vector), fp64.
 * rot / ty [n_cand, horizon] fp64 radians / metres: exactly the buffers a rollout plan reads (gm_rollout_desc.rot /
 * .ty), so a population's candidates go from the optimiser's vectors to rigid-particle targets without leaving the
 * device.  terms (optional) [n_cand, 3]: the curve regularisers of compute_loss -- ||vel / vel_div||, ||acc /
 * acc_div|| (compute_vel_acc / compute_vel_loss / compute_acc_loss, traj_utils.py:161-165, 230-246, 353-373) and
 * the penalty term (kind 0: compute_boundaries_penalty :263-274; kind 1: compute_vel_noninterp :418-436). */
enum gm_traj_kind { GM_TRAJ_INCREMENTS = 0, GM_TRAJ_PCHIP = 1 };
typedef struct gm_traj_param {
  int32_t kind;
  int32_t horizon;
  int32_t n_points;
  int32_t knot_stride;        /* kind 1 only */
  double rx_init;             /* radians */
  double scale_rot, scale_ty;
  double max_rot, max_ty;     /* radians, metres */
  double rotation_limit;      /* kind 0: penalty when the rotation leaves rx_init +- rotation_limit (2.8973) */
  double bound_penalty;       /* kind 0: 20.0 */
  double vel_div[2], acc_div[2];
} gm_traj_param;
int gm_traj_interpolate(const gm_traj_param* param_host, const double* x, int32_t n_cand, double* rot, double* ty,
                        double* terms, void* stream);

/* get_rigid_body_trajectory_from_diff / compute_particles_tmatrix (traj_utils.py:87-103, 167-194) for a whole
 * population: out [n_cand, steps, n_rigid, 3] fp32 rigid-particle targets from the control curves rot / ty
 * [n_cand, steps] (fp64) and the rigid particles' initial positions rigid_init [n_rigid, 3]. */
int gm_rigid_trajectory(const double* rot, const double* ty, int32_t n_cand, int32_t steps, const float* rigid_init,
                        int32_t n_rigid, const float* ty_init_host, float* out, void* stream);

/* ---- Row N1: batched Sinkhorn divergence (the Wasserstein term of compute_loss, traj_utils.py:276, 439) --------
 * loss[b] = S_eps(x_b, y) for uniform point clouds, p = 2, cost |x - y|^2 / 2, eps = blur^2, epsilon scaling with
 * ratio `scaling` from the clouds' joint diameter, symmetrised updates, debiased -- the tensorized algorithm of
 * geomloss.SamplesLoss(loss="sinkhorn", p=2, blur=.05) (traj_utils.py:69; geomloss defaults scaling=.5).  geomloss
 * is an unpinned third-party dependency absent from the reference checkout: PARITY UNPINNED (oracle: the fp64
 * numpy statement of the same algorithm, oracle/restate.py sinkhorn_divergence).
 *   x: point i of candidate b at x + b * x_batch_stride + (x_idx ? x_idx[i] : i) * x_stride, 3 floats -- e.g. the
 *      sand rows of the last frame of a rollout plan's obs, read in place
 *   y [m, 3]: target cloud shared by the population;  diameter_max: an upper bound of any candidate's joint
 *      bounding-box diagonal (sizes the fixed number of launched iterations; shorter schedules idle through) */
size_t gm_sinkhorn_workspace_bytes(int32_t n_batch, int32_t n, int32_t m);
int gm_sinkhorn_divergence(const float* x, int32_t x_stride, int64_t x_batch_stride, const int32_t* x_idx, int32_t n,
                           const float* y, int32_t m, int32_t n_batch, double blur, double scaling,
                           double diameter_max, float* loss, void* workspace, size_t workspace_bytes, void* stream);

/* ---- Rows E(2) + N3: training step building blocks (examples/train_dyn.py:49-72) -------------------------------
 * Every 128-wide Linear of forward, dgrad and wgrad is a tcgen05 GEMM (bf16x3 operand split, fp32 accumulation);
 * activations are fp32 [rows, 128].  gm_train_* below are the kernels gm_train_step is built from, exported so that
 * each one is parity-tested on its own (tests/test_gpu_train.py). */
/* pack n_jobs 128 x 128 blocks into 64 KB operand chunks; jobs_dev: device array of gm_pack_job */
typedef struct gm_pack_job {
  const float* w;       /* [128 rows][ld] fp32, device */
  int32_t ld, col0;     /* columns [col0, col0 + 128) */
  int32_t transpose;    /* 0: B = w (y = x w^T, forward);  1: B = w^T (dx = dy w, dgrad) */
  void* chunk;          /* 64 KB, 1024-byte aligned, device */
} gm_pack_job;
int gm_train_pack(const gm_pack_job* jobs_dev, int32_t n_jobs, void* stream);
/* out[r] (+)= mask o relu?( x[in_idx[r]] @ B^T + bias + add1[idx1[r]] + add2[idx2[r]] ); every optional pointer may be NULL */
int gm_train_gemm(const float* x, const int32_t* in_idx, const void* chunk, const float* bias, const float* add1,
                  const int32_t* idx1, const float* add2, const int32_t* idx2, const float* mask, int32_t relu,
                  int32_t accumulate, int32_t n_rows, const int32_t* n_rows_dev, float* out, void* stream);
/* dW[128, ld][:, col0:col0+128] += A^T B, dbias (optional) += column sums of A; partial: gm_train_wgrad_partial_bytes() */
size_t gm_train_wgrad_partial_bytes(void);
int gm_train_wgrad(const float* A, const float* B, int32_t n_rows, const int32_t* n_rows_dev, float* dW, int32_t ld,
                   int32_t col0, float* dbias, void* partial, void* stream);
int gm_train_layernorm_fwd(const float* z, const float* gamma, const float* beta, int32_t n_rows, float* xhat,
                           float* rstd, float* y, const float* res_in, float* res_out, void* stream);
size_t gm_train_layernorm_partial_bytes(int32_t n_rows);
int gm_train_layernorm_bwd(const float* d1, const float* d2, const int32_t* idx2, const float* xhat, const float* rstd,
                           const float* gamma, int32_t n_rows, float* dz, float* dgamma, float* dbeta, void* partial,
                           void* stream);
/* gm_train_layernorm_bwd without a scratch argument.  No counterpart in the reference: added
 * for callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
 * The per-block dgamma / dbeta partials live in a file-scope device array of the library (1 MB on every GPU that loads it,
 * until the process exits); the call allocates nothing.  Limit: n_rows <= GM_LN_STATIC_MAX_ROWS, larger sizes return
 * GM_STATUS_INVALID with a message naming the limit.  Same arithmetic and order as gm_train_layernorm_bwd (dgamma / dbeta
 * are accumulated into, dz is overwritten).  Every call shares the array: calls that overlap on the GPU (two streams, two
 * host threads) race on it; calls ordered on one stream do not.  Nothing in it needs clearing between calls. */
#define GM_LN_STATIC_MAX_ROWS 65536
int gm_train_layernorm_bwd_static(const float* d1, const float* d2, const int32_t* idx2, const float* xhat,
                                  const float* rstd, const float* gamma, int32_t n_rows, float* dz, float* dgamma,
                                  float* dbeta, void* stream);
int gm_train_small_fwd(const float* x, int32_t K, const float* w, const float* b, int32_t O, int32_t n_rows,
                       int32_t relu, float* h, void* stream);
size_t gm_train_small_partial_bytes(int32_t n_rows, int32_t K, int32_t O);
int gm_train_small_bwd(const float* x, int32_t K, const float* w, int32_t O, const float* dh, const float* mask,
                       int32_t n_rows, float* dW, float* db, float* dx, void* partial, void* stream);
/* L1Loss(reduction='sum') / n_selected_rows (train_dyn.py:56-64, 109); row_mask (bytes, optional) selects rows;
 * scratch: 2 + 2 * ceil(n_rows / 256) floats */
int gm_train_l1_loss(const float* pred, const float* tgt, const uint8_t* row_mask, int32_t n_rows, int32_t width,
                     float* loss, float* dpred, float* scratch, void* stream);
/* torch.optim.Adam step t = 1, 2, ... on a flat vector (train_dyn.py:98: Adam(lr), default betas / eps) */
int gm_adam_step(float* params, const float* grads, float* m, float* v, size_t n, float lr, float beta1, float beta2,
                 float eps, int32_t step, void* stream);

/* ---- the training step itself ---------------------------------------------------------------------------------
 * A plan owns the packed operand chunks of every 128 x 128 weight block (forward and transposed), the saved
 * activations of one batch and the backward scratch, all inside the caller's workspace.  params / grads: flat fp32
 * vectors in gm_model_create's parameter order (the optimiser updates params in place; call gm_train_sync_weights
 * afterwards).  Covers hidden = 128, num_layers = 2 (the reference configuration, rollout_utils.py:133-134). */
typedef struct gm_train gm_train;
typedef struct gm_train_desc {
  int32_t node_dim, edge_dim, out_dim, hidden, num_layers, m_steps;
  int32_t max_nodes, max_edges;     /* largest batch: nodes, edge slots (nodes * max_neighbours) */
  float* params;                    /* [gm_model_num_params] device */
  float* grads;                     /* same length, device: overwritten by every backward */
} gm_train_desc;
size_t gm_train_workspace_bytes(const gm_train_desc* desc_host);
int gm_train_create(gm_train** plan, const gm_train_desc* desc_host, void* workspace /*1024-byte aligned*/,
                    size_t workspace_bytes, void* stream);
void gm_train_destroy(gm_train* plan);
int gm_train_sync_weights(gm_train* plan, void* stream);
/* gradient buckets in the order the backward completes them (decoder, processor blocks m-1 .. 0, encoders): bucket b
 * is grads[offset, offset + count) */
int32_t gm_train_num_buckets(const gm_train* plan);
int gm_train_bucket(const gm_train* plan, int32_t bucket, size_t* offset_host, size_t* count_host);
const float* gm_train_prediction(const gm_train* plan);     /* [n_nodes, out_dim] of the last forward (device) */
/* test aid: the saved post-ReLU activations [rows, 128] of the last forward (edges in receiver-major order).
 * which: 0/1 edge encoder h1/h2, 2/3 node encoder h1/h2, 4/5 edge MLP h1/h2 of `step`, 6/7 node MLP h1/h2 of `step`,
 * 8/9 decoder h1/h2 */
const float* gm_train_activation(const gm_train* plan, int32_t which, int32_t step);
/* loss_batch (examples/train_dyn.py:49-72) without the optimiser step: forward on a batch of graphs given as one
 * disjoint graph (nodes [n, node_dim]; positions for the edge features; the receiver-major edge list of
 * gm_radius_graph: recv_ptr, esend, erecv, eref; edge_ptr = its sender-major offsets), loss = L1Loss(sum) / rows over
 * the rows row_mask selects (NULL = all), and -- with_backward -- every parameter gradient into grads.
 * grad_ready (optional, host callback) fires right after the kernels that complete bucket b were enqueued, so the
 * caller can order a collective on that slice behind them.  Everything is enqueued on `stream`. */
typedef void (*gm_grad_ready_fn)(void* user, int32_t bucket, size_t offset, size_t count);
int gm_train_forward_backward(gm_train* plan, const float* nodes, const float* pos, int32_t pos_stride, double conn_r,
                              int32_t n_nodes, int32_t max_edges, const int32_t* n_edges_dev, const int32_t* recv_ptr,
                              const int32_t* esend, const int32_t* erecv, const int32_t* eref, const int32_t* edge_ptr,
                              const float* target, const uint8_t* row_mask, float* loss, int32_t with_backward,
                              gm_grad_ready_fn grad_ready, void* user, void* stream);

/* ---- Row N3: training data path on the device -----------------------------------------------------------------
 * random_walk_noise (gnn_manip/utils/utils.py:96-115) + GraphBoundedMultimaterial._process_noisy
 * (collate_utils.py:169-193): per particle and cartesian axis, k-1 Normal(0, noise_std / sqrt(k-1)) velocity
 * increments (Philox4x32-10 keyed by seed, particle and axis), accumulated twice (velocity, then position); the
 * position noise is added to the cartesian columns of obs (first frame untouched) and its last frame to tgt.
 * samples (optional) [k-1, n, 3]: the raw normal draws, for tests. */
int gm_random_walk_noise(const float* obs, int32_t k, int32_t n, int32_t D, const gm_phys* phys, const float* tgt,
                         double noise_std, uint64_t seed, float* noisy_obs, float* noisy_tgt, float* samples,
                         void* stream);
/* compute_target (collate_utils.py:148-157): ((tgt - 2 p[-1] + p[-2]) - acc_mean) / acc_std -> [n, 3] */
int gm_training_target(const float* obs, int32_t k, int32_t n, int32_t D, const gm_phys* phys, const float* tgt,
                       float* target, void* stream);

/* Debug aid: phase timestamps (clock64) recorded by CTA 0 of the instrumented kernels; only in -DGM_STAMPS builds
 * (returns GM_STATUS_UNSUPPORTED otherwise).  out_host: host array of n <= 64 values. */
int gm_debug_stamps(unsigned long long* out_host, int32_t n);

/* tcgen05 operand-path self test: out[M,128] = x[M,128] @ w[128,128]^T (bf16x3), M % 128 == 0.
 * chunk: 64 KB device scratch for the packed weight. */
int gm_selftest_gemm(const float* x, const float* w, void* chunk, float* out, int32_t M, void* stream);
/* The same product through the tensor-memory operand path of the edge kernel (row bulk copies in, tcgen05.st
 * operand, tcgen05.mma with A in TMEM, bulk reduce-add out): out += x @ w^T, so zero `out` first. */
int gm_selftest_gemm_ts(const float* x, const float* w, void* chunk, float* out, int32_t M, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* GNN_MANIP_B200_H_ */
