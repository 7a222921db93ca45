// C ABI (include/gnn_manip_b200.h): model packing, forward orchestration, rollout plan + CUDA graph.
#include <math.h>
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>
#include <new>
#include <vector>

#include "../../include/gnn_manip_b200.h"
#include "common.cuh"
#include "mlp_simt.cuh"
#include "mp_tc.cuh"
#include "physics.cuh"
#include "radius_graph.cuh"
#include "scan.cuh"
#include "train.cuh"
#include "sinkhorn.cuh"
#include "traj.cuh"

namespace gm {

static thread_local char g_err[512] = "";
thread_local long long g_launch_count = 0;

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
int cuda_fail(cudaError_t e, const char* what) {
  set_error("CUDA error %d (%s) at %s", (int)e, cudaGetErrorString(e), what);
  return GM_ERR_CUDA;
}

thread_local Profiler* g_profiler = nullptr;
int Profiler::mark(int c) {
  if (n >= kMax) { set_error("profiler: more than %d marks", kMax); return GM_ERR_INVALID; }
  GM_CUDA(cudaEventCreate(&ev[n]));
  GM_CUDA(cudaEventRecord(ev[n], st));
  cls[n++] = c;
  return GM_OK;
}
int Profiler::finish(float* ms_by_phase, int* launches_by_phase) {
  cudaError_t ce = cudaStreamSynchronize(st);
  for (int c = 0; c < kProfCount; ++c) { ms_by_phase[c] = 0.f; launches_by_phase[c] = 0; }
  if (ce == cudaSuccess) {
    for (int i = 0; i + 1 < n; ++i) {
      float ms = 0.f;
      cudaEventElapsedTime(&ms, ev[i], ev[i + 1]);
      if (cls[i] >= 0 && cls[i] < kProfCount) { ms_by_phase[cls[i]] += ms; launches_by_phase[cls[i]] += 1; }
    }
  }
  for (int i = 0; i < n; ++i) cudaEventDestroy(ev[i]);
  n = 0;
  if (ce != cudaSuccess) return cuda_fail(ce, "profiler sync");
  return GM_OK;
}

namespace {

__global__ void transpose_kernel(const float* __restrict__ w, int rows, int cols, float* __restrict__ wt) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= rows * cols) return;
  const int r = i / cols, c = i - r * cols;
  wt[(size_t)c * rows + r] = w[i];
}

__global__ void edge_index_kernel(const int* __restrict__ nbr, const int* __restrict__ cnt,
                                  const int* __restrict__ edge_ptr, int n, int K, long long n_edges,
                                  long long* __restrict__ edge_index) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (long long)n * K) return;
  const int i = (int)(idx / K), s = (int)(idx - (long long)i * K);
  if (s >= cnt[i]) return;
  const long long e = (long long)edge_ptr[i] + s;
  if (e >= n_edges) return;
  edge_index[e] = i;
  edge_index[n_edges + e] = nbr[idx];
}

__global__ void next_position_kernel(const float* __restrict__ obs, int k, int n, int D, PhysParams p,
                                     const float* __restrict__ pred, float* __restrict__ next_pos) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n * 3) return;
  const int node = idx / 3, d = idx - 3 * node;
  const float acc = __fadd_rn(__fmul_rn(pred[idx], p.acc_std[d]), p.acc_mean[d]);
  const float lp = obs[((size_t)(k - 1) * n + node) * D + p.cart[d]];
  const float pp = obs[((size_t)(k - 2) * n + node) * D + p.cart[d]];
  next_pos[idx] = __fadd_rn(lp, __fadd_rn(__fsub_rn(lp, pp), acc));
}

__global__ void rigid_slot_kernel(const int* __restrict__ rigid_idx, int n_rigid, int* __restrict__ slot) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r < n_rigid) slot[rigid_idx[r]] = r;
}

PhysParams to_phys(const gm_phys& g, int k) {
  PhysParams p;
  memset(&p, 0, sizeof(p));
  for (int d = 0; d < 3; ++d) {
    p.cart[d] = g.cartesian_idx[d];
    p.ctrl[d] = g.control_idx[d];
    p.vel_mean[d] = g.velocity_mean[d];
    p.vel_std[d] = g.velocity_std[d];
    p.acc_mean[d] = g.acceleration_mean[d];
    p.acc_std[d] = g.acceleration_std[d];
    p.lower[d] = g.lower_bounds[d];
    p.upper[d] = g.upper_bounds[d];
  }
  p.mat = g.material_idx;
  p.has_ctrl = g.use_control ? 1 : 0;
  p.node_dim = 3 * (k - 1) + 6 + 1 + (p.has_ctrl ? 3 : 0);
  p.conn_r = (float)g.conn_r;
  return p;
}

struct MlpParams {
  int n_lin = 0;
  int in[kMaxSimtLayers], out[kMaxSimtLayers];
  const float* W[kMaxSimtLayers];
  const float* Wt[kMaxSimtLayers];
  const float* b[kMaxSimtLayers];
  const float* ln_g = nullptr;
  const float* ln_b = nullptr;
};

size_t mlp_param_count(int in1, int hidden, int out, int num_layers, bool norm) {
  size_t n = (size_t)hidden * in1 + hidden;
  for (int i = 0; i < num_layers - 1; ++i) n += (size_t)hidden * hidden + hidden;
  n += (size_t)out * hidden + out;
  if (norm) n += 2 * (size_t)out;
  return n;
}
size_t mlp_weight_count(int in1, int hidden, int out, int num_layers) {
  return (size_t)hidden * in1 + (size_t)(num_layers - 1) * hidden * hidden + (size_t)out * hidden;
}

}  // namespace
}  // namespace gm

using namespace gm;

#ifdef GM_TRACE
namespace gm {
int trace_dump_physics(unsigned long long*, int);
int trace_dump_radius(unsigned long long*, int);
int trace_dump_tc(unsigned long long*, int);
int trace_dump_edge(unsigned long long*, int);
int trace_dump_node16(unsigned long long*, int);
}
#endif

struct gm_model {
  int node_dim, edge_dim, out_dim, hidden, num_layers, m_steps;
  size_t n_params;
  MlpParams enc_edge, enc_node, dec;
  std::vector<MlpParams> proc_edge, proc_node;
  bool tc_ok;
  TcMlp tc_enc_edge, tc_enc_node, tc_dec;
  std::vector<TcMlp> tc_proc_edge, tc_proc_node;
};

namespace gm {
namespace {

bool shape_supports_tc(int node_dim, int edge_dim, int out_dim, int hidden, int num_layers) {
  return hidden == kH && num_layers == 2 && out_dim >= 1 && out_dim <= 3 && node_dim >= 1 && node_dim <= 32 &&
         edge_dim >= 1 && edge_dim <= 32;
}
size_t num_tc_chunks(int m_steps) { return 2 + 2 + 2 + (size_t)m_steps * 9; }

// carve one MLP's raw parameters (already copied into the blob) and its transposed weights
void carve_mlp(MlpParams& m, const float*& raw, float*& wt, int in1, int hidden, int out, int num_layers, bool norm) {
  m.n_lin = num_layers + 1;
  for (int l = 0; l < m.n_lin; ++l) {
    m.in[l] = l == 0 ? in1 : hidden;
    m.out[l] = l == m.n_lin - 1 ? out : hidden;
    m.W[l] = raw;
    raw += (size_t)m.in[l] * m.out[l];
    m.b[l] = raw;
    raw += m.out[l];
    m.Wt[l] = wt;
    wt += (size_t)m.in[l] * m.out[l];
  }
  if (norm) {
    m.ln_g = raw;
    raw += out;
    m.ln_b = raw;
    raw += out;
  }
}

int transpose_mlp(const MlpParams& m, cudaStream_t st) {
  for (int l = 0; l < m.n_lin; ++l) {
    const int tot = m.in[l] * m.out[l];
    transpose_kernel<<<ceil_div(tot, 256), 256, 0, st>>>(m.W[l], m.out[l], m.in[l], const_cast<float*>(m.Wt[l]));
  }
  GM_LAUNCH_CHECK("transpose_kernel");
  return GM_OK;
}

MlpSimtArgs simt_args(const MlpParams& m) {
  MlpSimtArgs a;
  memset(&a, 0, sizeof(a));
  a.n_layers = m.n_lin;
  for (int l = 0; l < m.n_lin; ++l) a.layers[l] = MlpSimtLayer{m.Wt[l], m.b[l], m.in[l], m.out[l]};
  a.ln_g = m.ln_g;
  a.ln_b = m.ln_b;
  a.ln_eps = 1e-5f;
  return a;
}

struct FwdWs {
  float *e, *v, *pr, *ps, *aggpart, *carry, *eo, *edge_attr_sorted;
};

void carve_forward(Arena& a, const gm_model* m, int n_nodes, int max_edges, bool simt, FwdWs& w) {
  const size_t H = m->hidden;
  w.e = a.take<float>((size_t)max_edges * H + 4);
  w.v = a.take<float>((size_t)n_nodes * H + 4);
  w.aggpart = a.take<float>((size_t)n_nodes * H + 4);
  if (simt) {
    w.eo = a.take<float>((size_t)max_edges * H + 4);
    w.edge_attr_sorted = a.take<float>((size_t)max_edges * m->edge_dim + 4);
    w.pr = w.ps = w.carry = nullptr;
  } else {
    w.pr = a.take<float>((size_t)n_nodes * H + 4);
    w.ps = a.take<float>((size_t)n_nodes * H + 4);
    w.carry = a.take<float>(((size_t)max_edges / kAggRows + 2) * H);
    w.eo = w.edge_attr_sorted = nullptr;
  }
}

bool use_tc(const gm_model* m, int impl) { return impl == GM_IMPL_TC || (impl == GM_IMPL_AUTO && m->tc_ok); }

int simt_encode(const gm_model* m, const float* nodes, const float* edge_attr, const float* pos, int pos_stride,
                float conn_r, int n_nodes, int max_edges, const int* n_edges_dev, const int* esend, const int* erecv,
                const int* eref, const FwdWs& w, cudaStream_t st) {
  int rc;
  // encoder: edges (receiver-major order) and nodes
  MlpSimtArgs a = simt_args(m->enc_edge);
  a.n_src = 1;
  if (edge_attr != nullptr) {
    a.src[0] = edge_attr; a.idx[0] = eref;
  } else {
    GM_REQUIRE(m->edge_dim == 4, "on-the-fly edge features need edge_dim == 4 (got %d)", m->edge_dim);
    if ((rc = edge_features(pos, pos_stride, esend, erecv, n_edges_dev, max_edges, conn_r, w.edge_attr_sorted, st)))
      return rc;
    a.src[0] = w.edge_attr_sorted; a.idx[0] = nullptr;
  }
  a.width[0] = m->edge_dim; a.src_stride[0] = m->edge_dim;
  a.M = max_edges; a.M_dev = n_edges_dev; a.out = w.e;
  if ((rc = mlp_simt(a, st))) return rc;
  a = simt_args(m->enc_node);
  a.n_src = 1; a.src[0] = nodes; a.width[0] = m->node_dim; a.src_stride[0] = m->node_dim;
  a.M = n_nodes; a.out = w.v;
  return mlp_simt(a, st);
}

int simt_mp_step(const gm_model* m, int s, int n_nodes, int max_edges, const int* n_edges_dev, const int* recv_ptr,
                 const int* esend, const int* erecv, const FwdWs& w, cudaStream_t st) {
  int rc;
  const int H = m->hidden;
  MlpSimtArgs a = simt_args(m->proc_edge[s]);          // phi_edge(cat[e, v[recv], v[send]])  (graph_networks.py:43-50)
  a.n_src = 3;
  a.src[0] = w.e; a.idx[0] = nullptr;
  a.src[1] = w.v; a.idx[1] = erecv;
  a.src[2] = w.v; a.idx[2] = esend;
  for (int i = 0; i < 3; ++i) { a.width[i] = H; a.src_stride[i] = H; }
  a.M = max_edges; a.M_dev = n_edges_dev;
  a.out_pre = w.eo; a.out = w.e; a.residual = w.e;
  if ((rc = mlp_simt(a, st))) return rc;
  if ((rc = segment_sum(w.eo, H, recv_ptr, nullptr, n_nodes, w.aggpart, st))) return rc;
  a = simt_args(m->proc_node[s]);          // phi_node(cat[agg, v])  (graph_networks.py:75-85)
  a.n_src = 2;
  a.src[0] = w.aggpart; a.src[1] = w.v;
  for (int i = 0; i < 2; ++i) { a.width[i] = H; a.src_stride[i] = H; a.idx[i] = nullptr; }
  a.M = n_nodes; a.out = w.v; a.residual = w.v;
  return mlp_simt(a, st);
}

int simt_decode(const gm_model* m, int n_nodes, float* pred, const FwdWs& w, cudaStream_t st) {
  const int H = m->hidden;
  MlpSimtArgs a = simt_args(m->dec);
  a.n_src = 1; a.src[0] = w.v; a.width[0] = H; a.src_stride[0] = H;
  a.M = n_nodes; a.out = pred;
  return mlp_simt(a, st);
}

int forward_simt(const gm_model* m, const float* nodes, const float* edge_attr, const float* pos, int pos_stride,
                 float conn_r, int n_nodes, int max_edges, const int* n_edges_dev, const int* recv_ptr,
                 const int* esend, const int* erecv, const int* eref, float* pred, const FwdWs& w,
                 cudaStream_t st) {
  int rc;
  if ((rc = simt_encode(m, nodes, edge_attr, pos, pos_stride, conn_r, n_nodes, max_edges, n_edges_dev, esend, erecv, eref,
                        w, st)))
    return rc;
  for (int s = 0; s < m->m_steps; ++s)
    if ((rc = simt_mp_step(m, s, n_nodes, max_edges, n_edges_dev, recv_ptr, esend, erecv, w, st))) return rc;
  return simt_decode(m, n_nodes, pred, w, st);
}

// node half of MP step s: the 16-warp kernel (mp_node16.cu); GM_NODE_KERNEL=8 selects the 8-warp kernel of mp_tc.cu (A/B aid)
int tc_node_half(const gm_model* m, int s, const TcMlp* next_edge, const TcMlp* dec, float* pred, const TcGraph& g,
                 const TcBuffers& b, cudaStream_t st) {
  static const bool node8 = [] { const char* e = getenv("GM_NODE_KERNEL"); return e && !strcmp(e, "8"); }();
  return node8 ? tc_mp_node(m->tc_proc_node[s], next_edge, dec, m->out_dim, pred, g, b, st)
               : tc_mp_node16(m->tc_proc_node[s], next_edge, dec, m->out_dim, pred, g, b, st);
}

int forward_tc(const gm_model* m, const float* nodes, const float* edge_attr, const float* pos, int pos_stride,
               float conn_r, int n_nodes, int max_edges, const int* n_edges_dev, const int* recv_ptr,
               const int* esend, const int* erecv, const int* eref, float* pred, const FwdWs& w,
               cudaStream_t st, cudaEvent_t enc_node_done = nullptr) {
  // enc_node_done: the caller already launched the node encoder on another stream (it needs only the node features,
  // so a rollout runs it beside the graph build); the first edge step joins on the event instead.
  GM_REQUIRE(m->tc_ok, "model shape (hidden=%d, num_layers=%d) is not covered by the tcgen05 kernels", m->hidden,
             m->num_layers);
  TcGraph g{n_edges_dev, max_edges, n_nodes, esend, erecv, recv_ptr};
  TcBuffers b{w.e, w.v, w.pr, w.ps, w.aggpart, w.carry};
  int rc;
  GM_PROF(kProfEncEdge);
  if ((rc = tc_enc_edge(m->tc_enc_edge, m->edge_dim, pos, pos_stride, conn_r, edge_attr, eref, g, b, st))) return rc;
  GM_PROF(kProfEncNode);
  if (enc_node_done == nullptr) {
    if ((rc = tc_enc_node(m->tc_enc_node, m->tc_proc_edge[0], m->node_dim, nodes, g, b, st))) return rc;
  } else {
    GM_CUDA(cudaStreamWaitEvent(st, enc_node_done, 0));
  }
  for (int s = 0; s < m->m_steps; ++s) {
    GM_PROF(kProfMpEdge);
    // after a join the first edge step is an ordinary launch: e then comes from the kernel right before it
    if ((rc = tc_mp_edge(m->tc_proc_edge[s], g, b, st, !(s == 0 && enc_node_done != nullptr)))) return rc;
    GM_PROF(kProfMpNode);
    const bool last = s + 1 == m->m_steps;
    rc = tc_node_half(m, s, last ? nullptr : &m->tc_proc_edge[s + 1], last ? &m->tc_dec : nullptr, pred, g, b, st);
    if (rc) return rc;
  }
  GM_PROF(-1);
  return GM_OK;
}

}  // namespace
}  // namespace gm

struct gm_rollout {
  gm_rollout_desc d;
  RolloutParams rp;
  RadiusGraphWs gws;
  FwdWs fws;
  int n, max_edges;
  int *nbr, *cnt, *edge_ptr, *recv_ptr, *esend, *erecv, *eref, *n_edges;
  float* nodes;
  float* pred;
  bool tc;
  cudaGraph_t graph = nullptr;
  cudaGraphExec_t exec = nullptr;
  int64_t launches_per_step = 0;
  cudaStream_t side = nullptr;            // node encoder runs here, beside the graph build
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
};

namespace gm {
namespace {

void carve_rollout(Arena& a, const gm_rollout_desc& d, gm_rollout& r) {
  const int n = d.n_per_graph * d.n_graphs, K = d.max_neighbours;
  const gm_model* m = d.model;
  r.n = n;
  r.max_edges = n * K;
  r.tc = use_tc(m, d.impl);
  radius_graph_carve(a, n, d.n_graphs, K, r.gws);
  r.nbr = a.take<int>((size_t)n * K);
  r.cnt = a.take<int>(n);
  r.edge_ptr = a.take<int>((size_t)n + 1);
  r.recv_ptr = a.take<int>((size_t)n + 1);
  r.esend = a.take<int>((size_t)n * K);
  r.erecv = a.take<int>((size_t)n * K);
  r.eref = a.take<int>((size_t)n * K);
  r.n_edges = a.take<int>(4);
  r.nodes = a.take<float>((size_t)n * m->node_dim + 4);
  r.pred = a.take<float>((size_t)n * m->out_dim + 4);
  r.rp.rigid_tgt = a.take<float>((size_t)d.n_graphs * (d.n_rigid > 0 ? d.n_rigid : 1) * 3);
  r.rp.rigid_slot = a.take<int>(d.n_per_graph);
  r.rp.step_dev = a.take<int>(4);
  carve_forward(a, m, n, r.max_edges, !r.tc, r.fws);
}

int rollout_step(gm_rollout* r, cudaStream_t st) {
  const gm_rollout_desc& d = r->d;
  const gm_model* m = d.model;
  const PhysParams& p = r->rp.phys;
  int rc;
  GM_PROF(kProfControl);
  if ((rc = apply_control(d.obs, r->rp, st))) return rc;
  const float* last = d.obs + (size_t)(d.k - 1) * r->n * d.D;
  const float* pos = last + p.cart[0];
  GM_PROF(kProfNodeFeat);
  if ((rc = node_features(d.obs, d.k, r->n, d.D, p, r->nodes, st))) return rc;
  const bool fork = r->tc && r->side != nullptr && g_profiler == nullptr;   // profiled steps run strictly serially
  if (fork) {
    // the node encoder needs the node features only: fork it onto the side stream so that it overlaps the
    // (many small) graph-build kernels; forward_tc joins before the first edge step
    TcGraph g{r->n_edges, r->max_edges, r->n, r->esend, r->erecv, r->recv_ptr};
    TcBuffers b{r->fws.e, r->fws.v, r->fws.pr, r->fws.ps, r->fws.aggpart, r->fws.carry};
    GM_CUDA(cudaEventRecord(r->ev_fork, st));
    GM_CUDA(cudaStreamWaitEvent(r->side, r->ev_fork, 0));
    if ((rc = tc_enc_node(m->tc_enc_node, m->tc_proc_edge[0], m->node_dim, r->nodes, g, b, r->side))) return rc;
    GM_CUDA(cudaEventRecord(r->ev_join, r->side));
  }
  if ((rc = radius_graph(pos, d.D, d.n_per_graph, d.n_graphs, d.phys.conn_r, d.max_neighbours, r->nbr, r->cnt,
                         r->edge_ptr, r->recv_ptr, r->esend, r->erecv, r->eref, r->n_edges, r->gws, st)))
    return rc;
  if (r->tc)
    rc = forward_tc(m, r->nodes, nullptr, pos, d.D, p.conn_r, r->n, r->max_edges, r->n_edges, r->recv_ptr, r->esend,
                    r->erecv, r->eref, r->pred, r->fws, st, fork ? r->ev_join : nullptr);
  else
    rc = forward_simt(m, r->nodes, nullptr, pos, d.D, p.conn_r, r->n, r->max_edges, r->n_edges, r->recv_ptr,
                      r->esend, r->erecv, r->eref, r->pred, r->fws, st);
  if (rc) return rc;
  GM_PROF(kProfIntegrate);
  if ((rc = integrate(d.obs, r->pred, r->rp, nullptr, st))) return rc;
  GM_PROF(-1);
  if (d.pred != nullptr) {
    GM_CUDA(cudaMemcpyAsync(d.pred, r->pred, sizeof(float) * (size_t)r->n * m->out_dim, cudaMemcpyDeviceToDevice, st));
  }
  return GM_OK;
}

}  // namespace
}  // namespace gm

// =====================================================================================================
extern "C" {

int gm_version(void) { return 100; }
long long gm_launch_count(void) { return gm::g_launch_count; }
const char* gm_last_error(void) { return gm::g_err; }

size_t gm_radius_graph_workspace_bytes(int32_t n_per_graph, int32_t n_graphs, int32_t K) {
  if (n_per_graph <= 0 || n_graphs <= 0 || K <= 0) return 0;
  return radius_graph_workspace_bytes(n_per_graph * n_graphs, n_graphs, K);
}

int gm_radius_graph(const float* pos, int32_t pos_stride, int32_t n_per_graph, int32_t n_graphs, double conn_r,
                    int32_t K, int32_t* nbr, int32_t* cnt, int32_t* edge_ptr, int32_t* recv_ptr, int32_t* esend,
                    int32_t* erecv, int32_t* eref, int32_t* n_edges, void* workspace, size_t workspace_bytes,
                    void* stream) {
  GM_REQUIRE(n_per_graph > 0 && n_graphs > 0 && K > 0, "gm_radius_graph: empty input");
  GM_REQUIRE(pos_stride >= 3, "gm_radius_graph: pos_stride=%d < 3", pos_stride);
  Arena a(workspace, workspace_bytes);
  RadiusGraphWs ws;
  radius_graph_carve(a, n_per_graph * n_graphs, n_graphs, K, ws);
  if (!a.ok || workspace == nullptr) {
    set_error("gm_radius_graph: workspace too small (%zu < %zu bytes)", workspace_bytes, a.off);
    return GM_ERR_WORKSPACE;
  }
  return radius_graph(pos, pos_stride, n_per_graph, n_graphs, conn_r, K, nbr, cnt, edge_ptr, recv_ptr, esend, erecv,
                      eref, n_edges, ws, (cudaStream_t)stream);
}

int gm_edge_index_i64(const int32_t* nbr, const int32_t* cnt, const int32_t* edge_ptr, int32_t n_nodes, int32_t K,
                      int64_t n_edges_host, int64_t* edge_index, void* stream) {
  if (n_nodes <= 0 || n_edges_host <= 0) return GM_OK;
  edge_index_kernel<<<ceil_div((long long)n_nodes * K, 256), 256, 0, (cudaStream_t)stream>>>(
      nbr, cnt, edge_ptr, n_nodes, K, n_edges_host, (long long*)edge_index);
  GM_LAUNCHED(1);
  GM_LAUNCH_CHECK("edge_index_kernel");
  return GM_OK;
}

size_t gm_recv_csr_workspace_bytes(int32_t n_nodes, int64_t n_edges) {
  Arena a(nullptr, 0);
  a.take<int>((size_t)n_nodes + 1);
  a.take<int>((size_t)n_nodes + 1);
  a.take<int>((size_t)(n_edges > 0 ? n_edges : 1));
  a.take<int>(scan_workspace_ints((long long)n_nodes + 1) + 8);
  return a.off + 256;
}

int gm_recv_csr(const int32_t* senders, const int32_t* receivers, int64_t n_edges, int32_t n_nodes,
                int32_t* recv_ptr, int32_t* esend, int32_t* erecv, int32_t* eref, void* workspace,
                size_t workspace_bytes, void* stream) {
  GM_REQUIRE(n_nodes > 0 && n_edges >= 0 && n_edges < (1LL << 31), "gm_recv_csr: bad sizes");
  Arena a(workspace, workspace_bytes);
  int* indeg = a.take<int>((size_t)n_nodes + 1);
  int* fill = a.take<int>((size_t)n_nodes + 1);
  int* keys = a.take<int>((size_t)(n_edges > 0 ? n_edges : 1));
  int* scan_ws = a.take<int>(scan_workspace_ints((long long)n_nodes + 1) + 8);
  if (!a.ok || workspace == nullptr) {
    set_error("gm_recv_csr: workspace too small (%zu < %zu bytes)", workspace_bytes, a.off);
    return GM_ERR_WORKSPACE;
  }
  return build_recv_csr(receivers, n_edges, nullptr, n_nodes, 0, nullptr, senders, indeg, fill, keys, scan_ws, false,
                        recv_ptr, esend, erecv, eref, (cudaStream_t)stream);
}

int32_t gm_node_dim(int32_t k, int32_t use_control) { return 3 * (k - 1) + 7 + (use_control ? 3 : 0); }

int gm_node_features(const float* obs, int32_t k, int32_t n, int32_t D, const gm_phys* phys, float* nodes,
                     void* stream) {
  GM_REQUIRE(phys != nullptr, "gm_node_features: phys is NULL");
  return node_features(obs, k, n, D, to_phys(*phys, k), nodes, (cudaStream_t)stream);
}

int gm_edge_features(const float* pos, int32_t pos_stride, const int32_t* senders, const int32_t* receivers,
                     const int32_t* n_edges_dev, int32_t max_edges, double conn_r, float* edge_attr, void* stream) {
  return edge_features(pos, pos_stride, senders, receivers, n_edges_dev, max_edges, (float)conn_r, edge_attr,
                       (cudaStream_t)stream);
}

size_t gm_model_num_params(int32_t node_dim, int32_t edge_dim, int32_t out_dim, int32_t hidden, int32_t num_layers,
                           int32_t m_steps) {
  return mlp_param_count(edge_dim, hidden, hidden, num_layers, true) +
         mlp_param_count(node_dim, hidden, hidden, num_layers, true) +
         (size_t)m_steps * (mlp_param_count(3 * hidden, hidden, hidden, num_layers, true) +
                            mlp_param_count(2 * hidden, hidden, hidden, num_layers, true)) +
         mlp_param_count(hidden, hidden, out_dim, num_layers, false);
}

static size_t model_weight_count(int node_dim, int edge_dim, int out_dim, int hidden, int num_layers, int m_steps) {
  return mlp_weight_count(edge_dim, hidden, hidden, num_layers) + mlp_weight_count(node_dim, hidden, hidden, num_layers) +
         (size_t)m_steps * (mlp_weight_count(3 * hidden, hidden, hidden, num_layers) +
                            mlp_weight_count(2 * hidden, hidden, hidden, num_layers)) +
         mlp_weight_count(hidden, hidden, out_dim, num_layers);
}

size_t gm_model_blob_bytes(int32_t node_dim, int32_t edge_dim, int32_t out_dim, int32_t hidden, int32_t num_layers,
                           int32_t m_steps) {
  size_t b = align_up(gm_model_num_params(node_dim, edge_dim, out_dim, hidden, num_layers, m_steps) * 4, 1024);
  b += align_up(model_weight_count(node_dim, edge_dim, out_dim, hidden, num_layers, m_steps) * 4, 1024);
  if (shape_supports_tc(node_dim, edge_dim, out_dim, hidden, num_layers)) b += num_tc_chunks(m_steps) * kChunkBytes;
  return b + 1024;
}

int gm_model_create(gm_model** out, const float* flat_params, size_t n_params, int32_t node_dim, int32_t edge_dim,
                    int32_t out_dim, int32_t hidden, int32_t num_layers, int32_t m_steps, void* blob,
                    size_t blob_bytes, void* stream) {
  GM_REQUIRE(out != nullptr && flat_params != nullptr && blob != nullptr, "gm_model_create: NULL argument");
  GM_REQUIRE(num_layers >= 2, "The number of layers num_layers must be at least 2");            // epd_gnn.py:26
  GM_REQUIRE(m_steps >= 1, "The number of m_steps message pasting steps must be at least 1");   // epd_gnn.py:27
  GM_REQUIRE(num_layers + 1 <= kMaxSimtLayers, "num_layers=%d exceeds the supported depth", num_layers);
  GM_REQUIRE(node_dim > 0 && edge_dim > 0 && out_dim > 0 && hidden > 0, "gm_model_create: non-positive dimension");
  const size_t expect = gm_model_num_params(node_dim, edge_dim, out_dim, hidden, num_layers, m_steps);
  GM_REQUIRE(n_params == expect, "gm_model_create: got %zu parameters, the architecture has %zu", n_params, expect);
  const size_t need = gm_model_blob_bytes(node_dim, edge_dim, out_dim, hidden, num_layers, m_steps);
  if (blob_bytes < need) {
    set_error("gm_model_create: blob too small (%zu < %zu bytes)", blob_bytes, need);
    return GM_ERR_WORKSPACE;
  }
  GM_REQUIRE(((uintptr_t)blob & 1023) == 0, "gm_model_create: blob must be 1024-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  {
    int rc0 = mlp_simt_configure();
    if (!rc0 && shape_supports_tc(node_dim, edge_dim, out_dim, hidden, num_layers)) rc0 = tc_configure();
    if (!rc0 && shape_supports_tc(node_dim, edge_dim, out_dim, hidden, num_layers)) rc0 = tc_configure_node16();
    if (rc0) return rc0;
  }
  gm_model* m = new (std::nothrow) gm_model();
  GM_REQUIRE(m != nullptr, "gm_model_create: out of host memory");
  m->node_dim = node_dim; m->edge_dim = edge_dim; m->out_dim = out_dim;
  m->hidden = hidden; m->num_layers = num_layers; m->m_steps = m_steps; m->n_params = n_params;
  m->tc_ok = shape_supports_tc(node_dim, edge_dim, out_dim, hidden, num_layers);

  char* base = (char*)blob;
  float* raw = (float*)base;
  cudaError_t ce = cudaMemcpyAsync(raw, flat_params, n_params * 4, cudaMemcpyDeviceToDevice, st);
  if (ce != cudaSuccess) { delete m; return cuda_fail(ce, "copy parameters"); }
  float* wt = (float*)(base + align_up(n_params * 4, 1024));
  char* chunks = (char*)wt + align_up(model_weight_count(node_dim, edge_dim, out_dim, hidden, num_layers, m_steps) * 4, 1024);

  const float* rp = raw;
  carve_mlp(m->enc_edge, rp, wt, edge_dim, hidden, hidden, num_layers, true);
  carve_mlp(m->enc_node, rp, wt, node_dim, hidden, hidden, num_layers, true);
  m->proc_edge.resize(m_steps);
  m->proc_node.resize(m_steps);
  for (int s = 0; s < m_steps; ++s) {
    carve_mlp(m->proc_edge[s], rp, wt, 3 * hidden, hidden, hidden, num_layers, true);
    carve_mlp(m->proc_node[s], rp, wt, 2 * hidden, hidden, hidden, num_layers, true);
  }
  carve_mlp(m->dec, rp, wt, hidden, hidden, out_dim, num_layers, false);

  int rc = transpose_mlp(m->enc_edge, st);
  if (!rc) rc = transpose_mlp(m->enc_node, st);
  for (int s = 0; s < m_steps && !rc; ++s) {
    rc = transpose_mlp(m->proc_edge[s], st);
    if (!rc) rc = transpose_mlp(m->proc_node[s], st);
  }
  if (!rc) rc = transpose_mlp(m->dec, st);

  if (!rc && m->tc_ok) {
    auto next_chunk = [&]() { char* c = chunks; chunks += kChunkBytes; return (void*)c; };
    auto base_tc = [](const MlpParams& p) {
      TcMlp t;
      memset(&t, 0, sizeof(t));
      t.b1 = p.b[0]; t.b2 = p.b[1]; t.b3 = p.b[2];
      t.ln_g = p.ln_g; t.ln_b = p.ln_b;
      return t;
    };
    auto pack = [&](const float* w, int ld, int col0, const void*& slot) {
      void* c = next_chunk();
      slot = c;
      if (!rc) rc = tc_pack_chunk(w, ld, col0, c, st);
    };
    // encoders: first layer on CUDA cores (fp32 [128][in]), layers 2,3 packed
    m->tc_enc_edge = base_tc(m->enc_edge);
    m->tc_enc_edge.w1 = m->enc_edge.W[0];
    pack(m->enc_edge.W[1], kH, 0, m->tc_enc_edge.wc[2]);
    pack(m->enc_edge.W[2], kH, 0, m->tc_enc_edge.wc[3]);
    m->tc_enc_node = base_tc(m->enc_node);
    m->tc_enc_node.w1 = m->enc_node.W[0];
    pack(m->enc_node.W[1], kH, 0, m->tc_enc_node.wc[2]);
    pack(m->enc_node.W[2], kH, 0, m->tc_enc_node.wc[3]);
    m->tc_proc_edge.resize(m_steps);
    m->tc_proc_node.resize(m_steps);
    for (int s = 0; s < m_steps; ++s) {
      TcMlp& e = m->tc_proc_edge[s];
      e = base_tc(m->proc_edge[s]);
      // W1 [128, 384]: columns 0:128 edge, 128:256 receiver, 256:384 sender (graph_networks.py:43-48)
      pack(m->proc_edge[s].W[0], 3 * kH, 0, e.wc[0]);
      pack(m->proc_edge[s].W[1], kH, 0, e.wc[1]);
      pack(m->proc_edge[s].W[2], kH, 0, e.wc[2]);
      pack(m->proc_edge[s].W[0], 3 * kH, kH, e.wc[3]);
      pack(m->proc_edge[s].W[0], 3 * kH, 2 * kH, e.wc[4]);
      TcMlp& nd = m->tc_proc_node[s];
      nd = base_tc(m->proc_node[s]);
      // W1 [128, 256]: columns 0:128 aggregated edges, 128:256 node (graph_networks.py:75-83)
      pack(m->proc_node[s].W[0], 2 * kH, 0, nd.wc[0]);
      pack(m->proc_node[s].W[0], 2 * kH, kH, nd.wc[1]);
      pack(m->proc_node[s].W[1], kH, 0, nd.wc[2]);
      pack(m->proc_node[s].W[2], kH, 0, nd.wc[3]);
    }
    m->tc_dec = base_tc(m->dec);
    pack(m->dec.W[0], kH, 0, m->tc_dec.wc[0]);
    pack(m->dec.W[1], kH, 0, m->tc_dec.wc[1]);
    m->tc_dec.w3 = m->dec.W[2];
  }
  if (rc) { delete m; return rc; }
  *out = m;
  return GM_OK;
}

void gm_model_destroy(gm_model* model) { delete model; }
int gm_model_supports_tc(const gm_model* model) { return model && model->tc_ok ? 1 : 0; }

size_t gm_forward_workspace_bytes(const gm_model* model, int32_t n_nodes, int32_t max_edges, int32_t impl) {
  if (!model || n_nodes <= 0 || max_edges < 0) return 0;
  Arena a(nullptr, 0);
  FwdWs w;
  carve_forward(a, model, n_nodes, max_edges, !use_tc(model, impl), w);
  return a.off + 256;
}

int gm_model_forward(const gm_model* model, int32_t impl, const float* nodes, const float* edge_attr,
                     const float* pos, int32_t pos_stride, double conn_r, int32_t n_nodes, int32_t max_edges,
                     const int32_t* n_edges_dev, const int32_t* recv_ptr, const int32_t* esend, const int32_t* erecv,
                     const int32_t* eref, float* pred, void* workspace, size_t workspace_bytes, void* stream) {
  GM_REQUIRE(model != nullptr && nodes != nullptr && pred != nullptr, "gm_model_forward: NULL argument");
  GM_REQUIRE(edge_attr != nullptr || pos != nullptr, "gm_model_forward: need edge_attr or pos");
  GM_REQUIRE(n_nodes > 0 && max_edges >= 0, "gm_model_forward: bad sizes");
  GM_REQUIRE(n_edges_dev != nullptr, "gm_model_forward: n_edges_dev is NULL");
  const bool tc = use_tc(model, impl);
  Arena a(workspace, workspace_bytes);
  FwdWs w;
  carve_forward(a, model, n_nodes, max_edges, !tc, w);
  if (!a.ok || workspace == nullptr) {
    set_error("gm_model_forward: workspace too small (%zu < %zu bytes)", workspace_bytes, a.off);
    return GM_ERR_WORKSPACE;
  }
  if (tc)
    return forward_tc(model, nodes, edge_attr, pos, pos_stride, (float)conn_r, n_nodes, max_edges, n_edges_dev,
                      recv_ptr, esend, erecv, eref, pred, w, (cudaStream_t)stream);
  return forward_simt(model, nodes, edge_attr, pos, pos_stride, (float)conn_r, n_nodes, max_edges, n_edges_dev,
                      recv_ptr, esend, erecv, eref, pred, w, (cudaStream_t)stream);
}

// ---- the three stages of EncProcDecGNN.forward as separate calls (epd_gnn.py:88-96); the latents live in the workspace
//      between the calls (same layout as gm_model_forward: gm_forward_workspace_bytes) ----
namespace {
int staged_ws(const char* who, const gm_model* model, int32_t impl, int32_t n_nodes, int32_t max_edges, void* workspace,
              size_t workspace_bytes, bool* tc, gm::FwdWs* w) {
  GM_REQUIRE(model != nullptr, "%s: model is NULL", who);
  GM_REQUIRE(n_nodes > 0 && max_edges >= 0, "%s: bad sizes", who);
  *tc = gm::use_tc(model, impl);
  GM_REQUIRE(!*tc || model->tc_ok, "%s: model shape (hidden=%d, num_layers=%d) is not covered by the tcgen05 kernels", who,
             model->hidden, model->num_layers);
  gm::Arena a(workspace, workspace_bytes);
  gm::carve_forward(a, model, n_nodes, max_edges, !*tc, *w);
  if (!a.ok || workspace == nullptr) {
    gm::set_error("%s: workspace too small (%zu < %zu bytes)", who, workspace_bytes, a.off);
    return gm::GM_ERR_WORKSPACE;
  }
  return gm::GM_OK;
}
}  // namespace

int gm_encode(const gm_model* model, int32_t impl, const float* nodes, const float* edge_attr, const float* pos,
              int32_t pos_stride, double conn_r, int32_t n_nodes, int32_t max_edges, const int32_t* n_edges_dev,
              const int32_t* recv_ptr, const int32_t* esend, const int32_t* erecv, const int32_t* eref, void* workspace,
              size_t workspace_bytes, void* stream) {
  GM_REQUIRE(nodes != nullptr && n_edges_dev != nullptr, "gm_encode: NULL argument");
  GM_REQUIRE(edge_attr != nullptr || pos != nullptr, "gm_encode: need edge_attr or pos");
  bool tc;
  FwdWs w;
  int rc = staged_ws("gm_encode", model, impl, n_nodes, max_edges, workspace, workspace_bytes, &tc, &w);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  if (!tc)
    return simt_encode(model, nodes, edge_attr, pos, pos_stride, (float)conn_r, n_nodes, max_edges, n_edges_dev, esend,
                       erecv, eref, w, st);
  TcGraph g{n_edges_dev, max_edges, n_nodes, esend, erecv, recv_ptr};
  TcBuffers b{w.e, w.v, w.pr, w.ps, w.aggpart, w.carry};
  if ((rc = tc_enc_edge(model->tc_enc_edge, model->edge_dim, pos, pos_stride, (float)conn_r, edge_attr, eref, g, b, st)))
    return rc;
  return tc_enc_node(model->tc_enc_node, model->tc_proc_edge[0], model->node_dim, nodes, g, b, st);
}

int gm_mp_step(const gm_model* model, int32_t impl, int32_t step, int32_t n_nodes, int32_t max_edges,
               const int32_t* n_edges_dev, const int32_t* recv_ptr, const int32_t* esend, const int32_t* erecv,
               void* workspace, size_t workspace_bytes, void* stream) {
  GM_REQUIRE(n_edges_dev != nullptr && recv_ptr != nullptr && esend != nullptr && erecv != nullptr,
             "gm_mp_step: NULL argument");
  bool tc;
  FwdWs w;
  int rc = staged_ws("gm_mp_step", model, impl, n_nodes, max_edges, workspace, workspace_bytes, &tc, &w);
  if (rc) return rc;
  GM_REQUIRE(step >= 0 && step < model->m_steps, "gm_mp_step: step=%d outside [0, %d)", step, model->m_steps);
  cudaStream_t st = (cudaStream_t)stream;
  if (!tc) return simt_mp_step(model, step, n_nodes, max_edges, n_edges_dev, recv_ptr, esend, erecv, w, st);
  TcGraph g{n_edges_dev, max_edges, n_nodes, esend, erecv, recv_ptr};
  TcBuffers b{w.e, w.v, w.pr, w.ps, w.aggpart, w.carry};
  // (an ordinary, fully ordered launch: the caller may have touched the latents between the stages)
  if ((rc = tc_mp_edge(model->tc_proc_edge[step], g, b, st, false))) return rc;
  // the node half also prepares Pr / Ps for the following edge step; after the last step there is none and the
  // (unused) projections are taken with this step's own edge MLP, the decoder being a stage of its own here
  const int nx = step + 1 < model->m_steps ? step + 1 : step;
  return tc_node_half(model, step, &model->tc_proc_edge[nx], nullptr, nullptr, g, b, st);
}

int gm_decode(const gm_model* model, int32_t impl, int32_t n_nodes, int32_t max_edges, float* pred, void* workspace,
              size_t workspace_bytes, void* stream) {
  GM_REQUIRE(pred != nullptr, "gm_decode: pred is NULL");
  bool tc;
  FwdWs w;
  int rc = staged_ws("gm_decode", model, impl, n_nodes, max_edges, workspace, workspace_bytes, &tc, &w);
  if (rc) return rc;
  // on both paths the decoder MLP runs on CUDA cores in fp32 (gm_model_forward fuses it into the last node kernel)
  return simt_decode(model, n_nodes, pred, w, (cudaStream_t)stream);
}

int gm_model_forward_profile(const gm_model* model, const float* nodes, const float* edge_attr, const float* pos,
                             int32_t pos_stride, double conn_r, int32_t n_nodes, int32_t max_edges,
                             const int32_t* n_edges_dev, const int32_t* recv_ptr, const int32_t* esend,
                             const int32_t* erecv, const int32_t* eref, float* pred, void* workspace,
                             size_t workspace_bytes, void* stream, float* ms_host, int32_t* launches_host) {
  GM_REQUIRE(model != nullptr && nodes != nullptr && pred != nullptr && ms_host != nullptr && launches_host != nullptr,
             "gm_model_forward_profile: NULL argument");
  GM_REQUIRE(model->tc_ok, "gm_model_forward_profile: the tcgen05 path does not cover this model shape");
  GM_REQUIRE(n_nodes > 0 && max_edges >= 0 && n_edges_dev != nullptr, "gm_model_forward_profile: bad sizes");
  Arena a(workspace, workspace_bytes);
  FwdWs w;
  carve_forward(a, model, n_nodes, max_edges, false, w);
  if (!a.ok || workspace == nullptr) {
    set_error("gm_model_forward_profile: workspace too small (%zu < %zu bytes)", workspace_bytes, a.off);
    return GM_ERR_WORKSPACE;
  }
  Profiler prof;
  prof.st = (cudaStream_t)stream;
  g_profiler = &prof;
  int rc = forward_tc(model, nodes, edge_attr, pos, pos_stride, (float)conn_r, n_nodes, max_edges, n_edges_dev,
                      recv_ptr, esend, erecv, eref, pred, w, prof.st);
  g_profiler = nullptr;
  float ms[kProfCount];
  int cnt[kProfCount];
  const int rc2 = prof.finish(ms, cnt);
  for (int c = 0; c < 4; ++c) { ms_host[c] = ms[kProfEncEdge + c]; launches_host[c] = cnt[kProfEncEdge + c]; }
  return rc ? rc : rc2;
}

int gm_mlp_forward(int32_t n_layers, const float* const* weights_t, const float* const* biases, const int32_t* dims,
                   const float* ln_weight, const float* ln_bias, int32_t n_src, const float* const* src,
                   const int32_t* const* idx, const int32_t* widths, int64_t n_rows, float* out, void* stream) {
  GM_REQUIRE(n_layers >= 1 && n_layers <= kMaxSimtLayers, "gm_mlp_forward: n_layers=%d", n_layers);
  GM_REQUIRE(n_src >= 1 && n_src <= 3, "gm_mlp_forward: n_src=%d", n_src);
  GM_REQUIRE(n_rows >= 0 && n_rows < (1LL << 31), "gm_mlp_forward: n_rows out of range");
  MlpSimtArgs a;
  memset(&a, 0, sizeof(a));
  a.n_layers = n_layers;
  for (int l = 0; l < n_layers; ++l) a.layers[l] = MlpSimtLayer{weights_t[l], biases ? biases[l] : nullptr, dims[l], dims[l + 1]};
  a.ln_g = ln_weight; a.ln_b = ln_bias; a.ln_eps = 1e-5f;
  GM_REQUIRE((ln_weight == nullptr) == (ln_bias == nullptr), "gm_mlp_forward: LayerNorm needs weight and bias");
  a.n_src = n_src;
  for (int s = 0; s < n_src; ++s) {
    a.src[s] = src[s]; a.idx[s] = idx ? idx[s] : nullptr; a.width[s] = widths[s]; a.src_stride[s] = widths[s];
  }
  a.M = (int)n_rows; a.out = out;
  return mlp_simt(a, (cudaStream_t)stream);
}

int gm_segment_sum(const float* src, int32_t width, const int32_t* recv_ptr, const int32_t* eref, int32_t n_nodes,
                   float* agg, void* stream) {
  return segment_sum(src, width, recv_ptr, eref, n_nodes, agg, (cudaStream_t)stream);
}

int gm_next_position(const float* obs, int32_t k, int32_t n, int32_t D, const gm_phys* phys, const float* pred,
                     float* next_pos, void* stream) {
  GM_REQUIRE(phys != nullptr && k >= 2 && n > 0, "gm_next_position: bad arguments");
  next_position_kernel<<<ceil_div(n * 3, 256), 256, 0, (cudaStream_t)stream>>>(obs, k, n, D, to_phys(*phys, k), pred,
                                                                              next_pos);
  GM_LAUNCHED(1);
  GM_LAUNCH_CHECK("next_position_kernel");
  return GM_OK;
}

size_t gm_rollout_workspace_bytes(const gm_rollout_desc* d) {
  if (!d || !d->model || d->n_per_graph <= 0 || d->n_graphs <= 0 || d->max_neighbours <= 0) return 0;
  Arena a(nullptr, 0);
  gm_rollout r;
  r.d = *d;
  carve_rollout(a, *d, r);
  return a.off + 256;
}

int gm_rollout_create(gm_rollout** plan, const gm_rollout_desc* d, void* workspace, size_t workspace_bytes,
                      void* stream) {
  GM_REQUIRE(plan != nullptr && d != nullptr && d->model != nullptr && d->obs != nullptr, "gm_rollout_create: NULL argument");
  GM_REQUIRE(d->n_per_graph > 0 && d->n_graphs > 0 && d->k >= 2 && d->D >= 3, "gm_rollout_create: bad shape");
  GM_REQUIRE(d->n_rigid >= 0 && d->n_rigid <= d->n_per_graph, "gm_rollout_create: n_rigid=%d", d->n_rigid);
  GM_REQUIRE(d->n_rigid == 0 || d->rigid_idx != nullptr, "gm_rollout_create: rigid_idx is NULL");
  GM_REQUIRE(d->n_rigid == 0 || d->traj_steps == 0 || d->traj != nullptr ||
                 (d->rot != nullptr && d->ty != nullptr && d->rigid_init != nullptr),
             "gm_rollout_create: need traj or (rot, ty, rigid_init)");
  GM_REQUIRE(d->n_rigid == 0 || d->phys.use_control, "gm_rollout_create: rigid control needs use_control");
  const gm_model* m = d->model;
  GM_REQUIRE(m->out_dim == 3, "rollout needs a 3-D acceleration model (out_dim=%d)", m->out_dim);
  GM_REQUIRE(m->edge_dim == 4, "rollout builds 4-D displacement edge features (edge_dim=%d)", m->edge_dim);
  GM_REQUIRE(m->node_dim == gm_node_dim(d->k, d->phys.use_control), "model node_dim=%d but features have %d columns",
             m->node_dim, gm_node_dim(d->k, d->phys.use_control));
  GM_REQUIRE((long long)d->n_per_graph * d->n_graphs * d->max_neighbours < (1LL << 31), "rollout: too many edge slots");
  gm_rollout* r = new (std::nothrow) gm_rollout();
  GM_REQUIRE(r != nullptr, "gm_rollout_create: out of host memory");
  r->d = *d;
  Arena a(workspace, workspace_bytes);
  carve_rollout(a, *d, *r);
  if (!a.ok || workspace == nullptr) {
    set_error("gm_rollout_create: workspace too small (%zu < %zu bytes)", workspace_bytes, a.off);
    delete r;
    return GM_ERR_WORKSPACE;
  }
  RolloutParams& rp = r->rp;
  rp.phys = to_phys(d->phys, d->k);
  rp.k = d->k; rp.D = d->D; rp.n_per_graph = d->n_per_graph; rp.n_graphs = d->n_graphs;
  rp.n_rigid = d->n_rigid; rp.rigid_idx = d->rigid_idx;
  rp.traj_steps = d->traj_steps; rp.traj = d->traj; rp.rot = d->rot; rp.ty = d->ty; rp.rigid_init = d->rigid_init;
  for (int i = 0; i < 3; ++i) rp.ty_init[i] = d->ty_init[i];
  cudaStream_t st = (cudaStream_t)stream;
  cudaError_t ce = cudaMemsetAsync(rp.step_dev, 0, 4 * sizeof(int), st);
  if (ce == cudaSuccess) ce = cudaMemsetAsync(const_cast<int*>(rp.rigid_slot), 0xFF, sizeof(int) * d->n_per_graph, st);
  if (ce != cudaSuccess) { delete r; return cuda_fail(ce, "rollout init"); }
  if (d->n_rigid > 0)
    rigid_slot_kernel<<<ceil_div(d->n_rigid, 128), 128, 0, st>>>(d->rigid_idx, d->n_rigid, const_cast<int*>(rp.rigid_slot));

  int rc = GM_OK;
  if (r->tc) {
    ce = cudaStreamCreateWithFlags(&r->side, cudaStreamNonBlocking);
    if (ce == cudaSuccess) ce = cudaEventCreateWithFlags(&r->ev_fork, cudaEventDisableTiming);
    if (ce == cudaSuccess) ce = cudaEventCreateWithFlags(&r->ev_join, cudaEventDisableTiming);
    if (ce != cudaSuccess) { gm_rollout_destroy(r); return cuda_fail(ce, "rollout side stream"); }
  }
  if (d->use_cuda_graph) {
    // capture one rollout step on a private stream; replays are ordered on the caller's stream
    cudaStream_t cs;
    ce = cudaStreamCreateWithFlags(&cs, cudaStreamNonBlocking);
    if (ce != cudaSuccess) { gm_rollout_destroy(r); return cuda_fail(ce, "cudaStreamCreate"); }
    ce = cudaStreamBeginCapture(cs, cudaStreamCaptureModeThreadLocal);
    if (ce == cudaSuccess) {
      const long long before = g_launch_count;
      rc = rollout_step(r, cs);
      r->launches_per_step = g_launch_count - before;
      cudaError_t ce2 = cudaStreamEndCapture(cs, &r->graph);
      if (rc == GM_OK && ce2 != cudaSuccess) rc = cuda_fail(ce2, "cudaStreamEndCapture");
      if (rc == GM_OK) {
        ce2 = cudaGraphInstantiate(&r->exec, r->graph, 0);
        if (ce2 != cudaSuccess) rc = cuda_fail(ce2, "cudaGraphInstantiate");
      }
    } else {
      rc = cuda_fail(ce, "cudaStreamBeginCapture");
    }
    cudaStreamDestroy(cs);
  }
  if (rc) { gm_rollout_destroy(r); return rc; }
  *plan = r;
  return GM_OK;
}

int gm_rollout_run(gm_rollout* r, int32_t steps, void* stream) {
  GM_REQUIRE(r != nullptr && steps >= 0, "gm_rollout_run: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  for (int i = 0; i < steps; ++i) {
    if (r->exec) {
      GM_CUDA(cudaGraphLaunch(r->exec, st));
    } else {
      const long long before = g_launch_count;
      int rc = rollout_step(r, st);
      if (rc) return rc;
      r->launches_per_step = g_launch_count - before;
    }
  }
  return GM_OK;
}

int gm_rollout_step_profile(gm_rollout* r, void* stream, float* ms_host, int32_t* launches_host) {
  GM_REQUIRE(r != nullptr && ms_host != nullptr && launches_host != nullptr, "gm_rollout_step_profile: NULL argument");
  Profiler prof;
  prof.st = (cudaStream_t)stream;
  g_profiler = &prof;
  const int rc = rollout_step(r, prof.st);
  g_profiler = nullptr;
  const int rc2 = prof.finish(ms_host, launches_host);
  return rc ? rc : rc2;
}

int gm_rollout_reset(gm_rollout* r, int32_t step, void* stream) {
  GM_REQUIRE(r != nullptr && step >= 0, "gm_rollout_reset: bad arguments");
  GM_CUDA(cudaMemcpyAsync(r->rp.step_dev, &step, sizeof(int), cudaMemcpyHostToDevice, (cudaStream_t)stream));
  GM_CUDA(cudaStreamSynchronize((cudaStream_t)stream));   // `step` lives on this call's stack
  return GM_OK;
}

int gm_rollout_graph(const gm_rollout* r, const int32_t** nbr, const int32_t** cnt, const int32_t** edge_ptr,
                     const int32_t** n_edges, const float** nodes) {
  GM_REQUIRE(r != nullptr, "gm_rollout_graph: NULL plan");
  if (nbr) *nbr = r->nbr;
  if (cnt) *cnt = r->cnt;
  if (edge_ptr) *edge_ptr = r->edge_ptr;
  if (n_edges) *n_edges = r->n_edges;
  if (nodes) *nodes = r->nodes;
  return GM_OK;
}

int64_t gm_rollout_kernel_launches_per_step(const gm_rollout* r) { return r ? r->launches_per_step : 0; }

void gm_rollout_destroy(gm_rollout* r) {
  if (!r) return;
  if (r->exec) cudaGraphExecDestroy(r->exec);
  if (r->graph) cudaGraphDestroy(r->graph);
  if (r->ev_fork) cudaEventDestroy(r->ev_fork);
  if (r->ev_join) cudaEventDestroy(r->ev_join);
  if (r->side) cudaStreamDestroy(r->side);
  delete r;
}

int gm_traj_interpolate(const gm_traj_param* q, const double* x, int32_t n_cand, double* rot, double* ty,
                        double* terms, void* stream) {
  GM_REQUIRE(q != nullptr && x != nullptr && rot != nullptr && ty != nullptr, "gm_traj_interpolate: NULL argument");
  TrajParams p;
  p.kind = q->kind; p.horizon = q->horizon; p.n_points = q->n_points; p.knot_stride = q->knot_stride;
  p.rx_init = q->rx_init; p.scale_rot = q->scale_rot; p.scale_ty = q->scale_ty;
  p.max_rot = q->max_rot; p.max_ty = q->max_ty;
  p.rotation_limit = q->rotation_limit; p.bound_penalty = q->bound_penalty;
  for (int i = 0; i < 2; ++i) { p.vel_div[i] = q->vel_div[i]; p.acc_div[i] = q->acc_div[i]; }
  return traj_interpolate(p, x, n_cand, rot, ty, terms, (cudaStream_t)stream);
}

int gm_rigid_trajectory(const double* rot, const double* ty, int32_t n_cand, int32_t steps, const float* rigid_init,
                        int32_t n_rigid, const float* ty_init_host, float* out, void* stream) {
  GM_REQUIRE(ty_init_host != nullptr, "gm_rigid_trajectory: ty_init is NULL");
  return rigid_trajectory(rot, ty, n_cand, steps, rigid_init, n_rigid, ty_init_host, out, (cudaStream_t)stream);
}

int gm_random_walk_noise(const float* obs, int32_t k, int32_t n, int32_t D, const gm_phys* phys, const float* tgt,
                         double noise_std, uint64_t seed, float* noisy_obs, float* noisy_tgt, float* samples,
                         void* stream) {
  GM_REQUIRE(phys != nullptr, "gm_random_walk_noise: phys is NULL");
  return random_walk_noise(obs, k, n, D, to_phys(*phys, k), tgt, noise_std, (unsigned long long)seed, noisy_obs,
                           noisy_tgt, samples, (cudaStream_t)stream);
}

int gm_training_target(const float* obs, int32_t k, int32_t n, int32_t D, const gm_phys* phys, const float* tgt,
                       float* target, void* stream) {
  GM_REQUIRE(phys != nullptr, "gm_training_target: phys is NULL");
  return training_target(obs, k, n, D, to_phys(*phys, k), tgt, target, (cudaStream_t)stream);
}

size_t gm_sinkhorn_workspace_bytes(int32_t n_batch, int32_t n, int32_t m) {
  if (n_batch <= 0 || n <= 0 || m <= 0) return 0;
  return sinkhorn_workspace_bytes(n_batch, n, m);
}

int gm_sinkhorn_divergence(const float* x, int32_t x_stride, int64_t x_batch_stride, const int32_t* x_idx, int32_t n,
                           const float* y, int32_t m, int32_t n_batch, double blur, double scaling,
                           double diameter_max, float* loss, void* workspace, size_t workspace_bytes, void* stream) {
  GM_REQUIRE(x != nullptr && y != nullptr && loss != nullptr, "gm_sinkhorn_divergence: NULL argument");
  GM_REQUIRE(n > 0 && m > 0 && n_batch > 0 && x_stride >= 3, "gm_sinkhorn_divergence: bad sizes");
  GM_REQUIRE(blur > 0.0 && scaling > 0.0 && scaling < 1.0 && diameter_max > 0.0,
             "gm_sinkhorn_divergence: blur=%g scaling=%g diameter_max=%g", blur, scaling, diameter_max);
  Arena a(workspace, workspace_bytes);
  SinkhornWs ws;
  sinkhorn_carve(a, n_batch, n, m, ws);
  if (!a.ok || workspace == nullptr) {
    set_error("gm_sinkhorn_divergence: workspace too small (%zu < %zu bytes)", workspace_bytes, a.off);
    return GM_ERR_WORKSPACE;
  }
  // schedule length for the largest diameter: [d^2] + arange(2 log d, 2 log blur, 2 log scaling) + [blur^2]
  int iters = 2;
  if (diameter_max > blur) iters += (int)ceil(log(diameter_max / blur) / -log(scaling));
  if (iters > kSinkMaxEps) iters = kSinkMaxEps;
  return sinkhorn_divergence(x, x_stride, x_batch_stride, x_idx, n, y, m, n_batch, blur, scaling, iters, loss, ws,
                             (cudaStream_t)stream);
}

static_assert(sizeof(gm_pack_job) == sizeof(TrPackJob), "gm_pack_job and TrPackJob must have the same layout");

int gm_train_pack(const gm_pack_job* jobs_dev, int32_t n_jobs, void* stream) {
  return tr_pack(reinterpret_cast<const TrPackJob*>(jobs_dev), n_jobs, (cudaStream_t)stream);
}

int gm_train_gemm(const float* x, const int32_t* in_idx, const void* chunk, const float* bias, const float* add1,
                  const int32_t* idx1, const float* add2, const int32_t* idx2, const float* mask, int32_t relu,
                  int32_t accumulate, int32_t n_rows, const int32_t* n_rows_dev, float* out, void* stream) {
  TrGemmArgs a{x, in_idx, chunk, bias, add1, idx1, add2, idx2, mask, out, relu, accumulate, n_rows, n_rows_dev};
  return tr_gemm(a, (cudaStream_t)stream);
}

size_t gm_train_wgrad_partial_bytes(void) { return tr_wgrad_partial_floats() * sizeof(float); }

int gm_train_wgrad(const float* A, const float* B, int32_t n_rows, const int32_t* n_rows_dev, float* dW, int32_t ld,
                   int32_t col0, float* dbias, void* partial, void* stream) {
  return tr_wgrad(A, B, n_rows, n_rows_dev, dW, ld, col0, dbias, (float*)partial, (cudaStream_t)stream);
}

int gm_train_layernorm_fwd(const float* z, const float* gamma, const float* beta, int32_t n_rows, float* xhat,
                           float* rstd, float* y, const float* res_in, float* res_out, void* stream) {
  return tr_ln_fwd(z, gamma, beta, n_rows, nullptr, xhat, rstd, y, res_in, res_out, (cudaStream_t)stream);
}

size_t gm_train_layernorm_partial_bytes(int32_t n_rows) { return tr_ln_partial_floats(n_rows) * sizeof(float); }

int gm_train_layernorm_bwd(const float* d1, const float* d2, const int32_t* idx2, const float* xhat, const float* rstd,
                           const float* gamma, int32_t n_rows, float* dz, float* dgamma, float* dbeta, void* partial,
                           void* stream) {
  return tr_ln_bwd(d1, d2, idx2, xhat, rstd, gamma, n_rows, nullptr, dz, dgamma, dbeta, (float*)partial,
                   (cudaStream_t)stream);
}

int gm_train_layernorm_bwd_static(const float* d1, const float* d2, const int32_t* idx2, const float* xhat,
                                  const float* rstd, const float* gamma, int32_t n_rows, float* dz, float* dgamma,
                                  float* dbeta, void* stream) {
  return tr_ln_bwd_static(d1, d2, idx2, xhat, rstd, gamma, n_rows, dz, dgamma, dbeta, (cudaStream_t)stream);
}

int gm_train_small_fwd(const float* x, int32_t K, const float* w, const float* b, int32_t O, int32_t n_rows,
                       int32_t relu, float* h, void* stream) {
  return tr_small_fwd(x, K, w, b, O, n_rows, nullptr, relu, h, (cudaStream_t)stream);
}

size_t gm_train_small_partial_bytes(int32_t n_rows, int32_t K, int32_t O) {
  return tr_small_partial_floats(n_rows, K, O) * sizeof(float);
}

int gm_train_small_bwd(const float* x, int32_t K, const float* w, int32_t O, const float* dh, const float* mask,
                       int32_t n_rows, float* dW, float* db, float* dx, void* partial, void* stream) {
  return tr_small_bwd(x, K, w, O, dh, mask, n_rows, nullptr, dW, db, dx, nullptr, (float*)partial, (cudaStream_t)stream);
}

int gm_train_l1_loss(const float* pred, const float* tgt, const uint8_t* row_mask, int32_t n_rows, int32_t width,
                     float* loss, float* dpred, float* scratch, void* stream) {
  return tr_l1_loss(pred, tgt, row_mask, n_rows, width, loss, dpred, scratch, (cudaStream_t)stream);
}

int gm_adam_step(float* params, const float* grads, float* m, float* v, size_t n, float lr, float beta1, float beta2,
                 float eps, int32_t step, void* stream) {
  return tr_adam(params, grads, m, v, n, lr, beta1, beta2, eps, step, (cudaStream_t)stream);
}

#ifdef GM_TRACE
// -DGM_TRACE: four sections of 1 + 160 entries (count, then the %globaltimer stamps): physics, radius_graph, mp_tc, mp_edge
int gm_debug_stamps(unsigned long long* out_host, int32_t n) {
  if (n < 4 * 161) return GM_ERR_UNSUPPORTED;
  cudaDeviceSynchronize();
  int (*fn[5])(unsigned long long*, int) = {gm::trace_dump_physics, gm::trace_dump_radius, gm::trace_dump_tc, gm::trace_dump_edge,
                                            gm::trace_dump_node16};
  const int nf = n >= 5 * 161 ? 5 : 4;          // (a fifth block, mp_node16.cu, for callers that pass room for it)
  for (int i = 0; i < nf; ++i) out_host[i * 161] = (unsigned long long)fn[i](out_host + i * 161 + 1, 160);
  return GM_OK;
}
#else
int gm_debug_stamps(unsigned long long* out_host, int32_t n) { return tc_debug_stamps(out_host, n); }
#endif

int gm_selftest_gemm(const float* x, const float* w, void* chunk, float* out, int32_t M, void* stream) {
  int rc = tc_configure();
  if (rc) return rc;
  rc = tc_pack_chunk(w, kH, 0, chunk, (cudaStream_t)stream);
  if (rc) return rc;
  return tc_selftest_gemm(x, chunk, out, M, (cudaStream_t)stream);
}

int gm_selftest_gemm_ts(const float* x, const float* w, void* chunk, float* out, int32_t M, void* stream) {
  int rc = tc_pack_chunk(w, kH, 0, chunk, (cudaStream_t)stream);
  if (rc) return rc;
  return tc_selftest_gemm_ts(x, chunk, out, M, (cudaStream_t)stream);
}

}  // extern "C"
