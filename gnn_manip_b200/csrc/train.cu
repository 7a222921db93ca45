// Rows E(2) + N3 -- one training step of EncProcDecGNN on a batch of graphs: forward with saved activations, L1 loss,
// backward (examples/train_dyn.py:49-72 `loss_batch`; the model is gnn_manip/models/epd_gnn.py:86-105 over
// deps/torch-graphnet/torch_graphnet/graph_networks.py:29-50, 63-85).  Built from the kernels of train_kernels.cu.
//
// Layout: edges live in receiver-major order (rows of a receiver contiguous, recv_ptr), as everywhere in this
// library; the sender-side sums of the backward use the reference's own sender-major edge ids (edge_ptr + the inverse of
// eref).  The first edge layer is split algebraically as in the inference kernels,
//   [e | v_recv | v_send] W1^T = e We^T + (v Wr^T + b1)[recv] + (v Ws^T)[send],
// so that in the backward the receiver / sender parts of dW1 and of the node gradient are NODE-level GEMMs over the
// segment sums of dz1 (20 x fewer rows than edges), and only e We^T costs a per-edge wgrad / dgrad.
// Saved per message-passing step: e_s, h1, h2, xhat, rstd on edges; v_s, agg_s, h1, h2, xhat, rstd on nodes (fp32).
// Gradients land in one flat fp32 vector in the parameter order of gm_model_create (bucket per processor block, so a
// caller may all-reduce block s while the backward of block s - 1 runs); weight-gradient reductions are deterministic.
#include <string.h>
#include <new>
#include <vector>

#include "../../include/gnn_manip_b200.h"
#include "mlp_simt.cuh"
#include "physics.cuh"
#include "train.cuh"

namespace gm {
namespace {

constexpr int H = kTrH;

struct MlpRef {            // offsets (floats) into the flat parameter / gradient vectors
  size_t W1, b1, W2, b2, W3, b3, g, b;
  int in1, out3;
  bool norm;
};

// one 128 x 128 block of a weight matrix and its two operand chunks
struct Block {
  size_t w_off;            // offset of the matrix the block lives in
  int ld, col0;
  void* fwd;               // B = W   (y = x W^T)
  void* bwd;               // B = W^T (dx = dy W)
};

struct EdgeMlp { MlpRef p; Block We, Wr, Ws, W2, W3; };
struct NodeMlp { MlpRef p; Block W1a, W1b, W2, W3; };
struct EncMlp { MlpRef p; Block W2, W3; };
struct DecMlp { MlpRef p; Block W1, W2; };

size_t carve_mlp_ref(MlpRef& r, size_t off, int in1, int out3, bool norm) {
  r.in1 = in1; r.out3 = out3; r.norm = norm;
  r.W1 = off; off += (size_t)H * in1;
  r.b1 = off; off += H;
  r.W2 = off; off += (size_t)H * H;
  r.b2 = off; off += H;
  r.W3 = off; off += (size_t)out3 * H;
  r.b3 = off; off += out3;
  r.g = r.b = 0;
  if (norm) { r.g = off; off += out3; r.b = off; off += out3; }
  return off;
}

struct StepAct {           // saved activations of one message-passing step
  float *e_in, *eh1, *eh2, *exhat, *erstd;
  float *v_in, *agg, *nh1, *nh2, *nxhat, *nrstd;
};

}  // namespace
}  // namespace gm

using namespace gm;

struct gm_train {
  gm_train_desc d;
  size_t n_params;
  EncMlp enc_edge, enc_node;
  std::vector<EdgeMlp> pe;
  std::vector<NodeMlp> pn;
  DecMlp dec;
  std::vector<Block*> blocks;
  TrPackJob* jobs_dev;
  int n_jobs;
  // activations
  float *ea, *ee_h1, *ee_h2, *ee_xhat, *ee_rstd;
  float *en_h1, *en_h2, *en_xhat, *en_rstd;
  std::vector<StepAct> act;
  float *e_last, *v_last;            // e_m (unused downstream), v_m (decoder input)
  float *dc_h1, *dc_h2, *pred, *dpred;
  // scratch
  float *zE, *yE, *zN, *tN, *Pr, *Ps;
  float *de, *dzE_a, *dzE_b, *dv, *dzN_a, *dzN_b, *dagg, *dPr, *dPs;
  float *wg_partial, *ln_partial, *small_partial, *loss_scratch;
  int* inv_eref;
  size_t bucket_off[16], bucket_cnt[16];
  int n_buckets;
};

namespace gm {
namespace {

__global__ void invert_perm_kernel(const int* __restrict__ eref, int max_edges, const int* __restrict__ n_dev,
                                   int* __restrict__ inv) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int n = n_dev ? min(*n_dev, max_edges) : max_edges;
  if (i < n) inv[eref[i]] = i;
}

void add_block(gm_train& t, Block& b, size_t w_off, int ld, int col0) {
  b.w_off = w_off; b.ld = ld; b.col0 = col0; b.fwd = b.bwd = nullptr;
  t.blocks.push_back(&b);
}

// everything the plan needs from the caller's workspace; base == nullptr only measures
void carve_train(Arena& a, gm_train& t) {
  const gm_train_desc& d = t.d;
  const size_t E = (size_t)d.max_edges, N = (size_t)d.max_nodes;
  const int m = d.m_steps;
  for (Block* b : t.blocks) {
    a.off = align_up(a.off, 1024);
    b->fwd = a.take<char>(kTrChunkBytes);
    a.off = align_up(a.off, 1024);
    b->bwd = a.take<char>(kTrChunkBytes);
  }
  t.jobs_dev = a.take<TrPackJob>(2 * t.blocks.size());
  auto edge = [&]() { return a.take<float>(E * H + 4); };
  auto node = [&]() { return a.take<float>(N * H + 4); };
  t.ea = a.take<float>(E * d.edge_dim + 4);
  t.ee_h1 = edge(); t.ee_h2 = edge(); t.ee_xhat = edge(); t.ee_rstd = a.take<float>(E + 4);
  t.en_h1 = node(); t.en_h2 = node(); t.en_xhat = node(); t.en_rstd = a.take<float>(N + 4);
  t.act.resize(m);
  for (int s = 0; s < m; ++s) {
    StepAct& c = t.act[s];
    c.e_in = edge(); c.eh1 = edge(); c.eh2 = edge(); c.exhat = edge(); c.erstd = a.take<float>(E + 4);
    c.v_in = node(); c.agg = node(); c.nh1 = node(); c.nh2 = node(); c.nxhat = node(); c.nrstd = a.take<float>(N + 4);
  }
  t.e_last = edge(); t.v_last = node();
  t.dc_h1 = node(); t.dc_h2 = node();
  t.pred = a.take<float>(N * d.out_dim + 4); t.dpred = a.take<float>(N * d.out_dim + 4);
  t.zE = edge(); t.yE = edge(); t.zN = node(); t.tN = node(); t.Pr = node(); t.Ps = node();
  t.de = edge(); t.dzE_a = edge(); t.dzE_b = edge();
  t.dv = node(); t.dzN_a = node(); t.dzN_b = node(); t.dagg = node(); t.dPr = node(); t.dPs = node();
  t.wg_partial = a.take<float>(tr_wgrad_partial_floats());
  t.ln_partial = a.take<float>(tr_ln_partial_floats((int)(E > N ? E : N)));
  size_t sp = tr_small_partial_floats((int)E, d.edge_dim, H);
  const size_t sp2 = tr_small_partial_floats((int)N, d.node_dim, H), sp3 = tr_small_partial_floats((int)N, H, d.out_dim);
  sp = sp > sp2 ? sp : sp2;
  sp = sp > sp3 ? sp : sp3;
  t.small_partial = a.take<float>(sp);
  t.loss_scratch = a.take<float>(2 + 2 * (size_t)ceil_div((long long)N, 256) + 4);
  t.inv_eref = a.take<int>(E + 4);
}

void layout_model(gm_train& t) {
  const gm_train_desc& d = t.d;
  size_t off = 0;
  t.blocks.clear();
  off = carve_mlp_ref(t.enc_edge.p, off, d.edge_dim, H, true);
  off = carve_mlp_ref(t.enc_node.p, off, d.node_dim, H, true);
  t.pe.resize(d.m_steps);
  t.pn.resize(d.m_steps);
  for (int s = 0; s < d.m_steps; ++s) {
    off = carve_mlp_ref(t.pe[s].p, off, 3 * H, H, true);
    off = carve_mlp_ref(t.pn[s].p, off, 2 * H, H, true);
  }
  off = carve_mlp_ref(t.dec.p, off, H, d.out_dim, false);
  t.n_params = off;
  add_block(t, t.enc_edge.W2, t.enc_edge.p.W2, H, 0); add_block(t, t.enc_edge.W3, t.enc_edge.p.W3, H, 0);
  add_block(t, t.enc_node.W2, t.enc_node.p.W2, H, 0); add_block(t, t.enc_node.W3, t.enc_node.p.W3, H, 0);
  for (int s = 0; s < d.m_steps; ++s) {
    EdgeMlp& e = t.pe[s];
    add_block(t, e.We, e.p.W1, 3 * H, 0); add_block(t, e.Wr, e.p.W1, 3 * H, H); add_block(t, e.Ws, e.p.W1, 3 * H, 2 * H);
    add_block(t, e.W2, e.p.W2, H, 0); add_block(t, e.W3, e.p.W3, H, 0);
    NodeMlp& n = t.pn[s];
    add_block(t, n.W1a, n.p.W1, 2 * H, 0); add_block(t, n.W1b, n.p.W1, 2 * H, H);
    add_block(t, n.W2, n.p.W2, H, 0); add_block(t, n.W3, n.p.W3, H, 0);
  }
  add_block(t, t.dec.W1, t.dec.p.W1, H, 0); add_block(t, t.dec.W2, t.dec.p.W2, H, 0);
  // gradient buckets in the order the backward finishes them: decoder, processor blocks m-1 .. 0, encoders
  int nb = 0;
  t.bucket_off[nb] = t.dec.p.W1; t.bucket_cnt[nb++] = t.n_params - t.dec.p.W1;
  for (int s = d.m_steps - 1; s >= 0 && nb < 15; --s) {
    const size_t end = s + 1 < d.m_steps ? t.pe[s + 1].p.W1 : t.dec.p.W1;
    t.bucket_off[nb] = t.pe[s].p.W1; t.bucket_cnt[nb++] = end - t.pe[s].p.W1;
  }
  t.bucket_off[nb] = 0; t.bucket_cnt[nb++] = d.m_steps > 0 ? t.pe[0].p.W1 : t.dec.p.W1;
  t.n_buckets = nb;
}

bool desc_ok(const gm_train_desc* d) {
  return d && d->hidden == H && d->num_layers == 2 && d->m_steps >= 1 && d->m_steps <= 14 && d->node_dim >= 1 &&
         d->node_dim <= 32 && d->edge_dim >= 1 && d->edge_dim <= 32 && d->out_dim >= 1 && d->out_dim <= 8 &&
         d->max_nodes > 0 && d->max_edges > 0;
}

}  // namespace
}  // namespace gm

extern "C" {

size_t gm_train_workspace_bytes(const gm_train_desc* d) {
  if (!desc_ok(d)) return 0;
  gm_train t;
  t.d = *d;
  layout_model(t);
  Arena a(nullptr, 0);
  carve_train(a, t);
  return a.off + 2048;
}

int gm_train_create(gm_train** plan, const gm_train_desc* d, void* workspace, size_t workspace_bytes, void* stream) {
  GM_REQUIRE(plan != nullptr && d != nullptr && workspace != nullptr, "gm_train_create: NULL argument");
  GM_REQUIRE(desc_ok(d), "gm_train_create: the training kernels cover hidden=128, num_layers=2, m_steps<=14, "
             "node_dim/edge_dim<=32, out_dim<=8 (got hidden=%d num_layers=%d m_steps=%d)", d->hidden, d->num_layers, d->m_steps);
  GM_REQUIRE(d->params != nullptr && d->grads != nullptr, "gm_train_create: params / grads are NULL");
  gm_train* t = new (std::nothrow) gm_train();
  GM_REQUIRE(t != nullptr, "gm_train_create: out of host memory");
  t->d = *d;
  layout_model(*t);
  const size_t expect = gm_model_num_params(d->node_dim, d->edge_dim, d->out_dim, d->hidden, d->num_layers, d->m_steps);
  if (t->n_params != expect) {
    set_error("gm_train_create: parameter layout mismatch (%zu vs %zu)", t->n_params, expect);
    delete t;
    return GM_ERR_INVALID;
  }
  if (((uintptr_t)workspace & 1023) != 0) {
    set_error("gm_train_create: workspace must be 1024-byte aligned");
    delete t;
    return GM_ERR_INVALID;
  }
  Arena a(workspace, workspace_bytes);
  carve_train(a, *t);
  if (!a.ok) {
    set_error("gm_train_create: workspace too small (%zu < %zu bytes)", workspace_bytes, a.off);
    delete t;
    return GM_ERR_WORKSPACE;
  }
  std::vector<TrPackJob> jobs;
  for (Block* b : t->blocks) {
    jobs.push_back(TrPackJob{d->params + b->w_off, b->ld, b->col0, 0, b->fwd});
    jobs.push_back(TrPackJob{d->params + b->w_off, b->ld, b->col0, 1, b->bwd});
  }
  t->n_jobs = (int)jobs.size();
  cudaError_t ce = cudaMemcpyAsync(t->jobs_dev, jobs.data(), sizeof(TrPackJob) * jobs.size(), cudaMemcpyHostToDevice,
                                   (cudaStream_t)stream);
  if (ce == cudaSuccess) ce = cudaStreamSynchronize((cudaStream_t)stream);     // `jobs` lives on this stack
  if (ce != cudaSuccess) { delete t; return cuda_fail(ce, "gm_train_create: upload pack jobs"); }
  int rc = gm_train_sync_weights(t, stream);
  if (rc) { delete t; return rc; }
  *plan = t;
  return GM_OK;
}

void gm_train_destroy(gm_train* plan) { delete plan; }

int gm_train_sync_weights(gm_train* t, void* stream) {
  GM_REQUIRE(t != nullptr, "gm_train_sync_weights: NULL plan");
  return tr_pack(t->jobs_dev, t->n_jobs, (cudaStream_t)stream);
}

int32_t gm_train_num_buckets(const gm_train* t) { return t ? t->n_buckets : 0; }

int gm_train_bucket(const gm_train* t, int32_t bucket, size_t* offset, size_t* count) {
  GM_REQUIRE(t != nullptr && bucket >= 0 && bucket < t->n_buckets && offset && count, "gm_train_bucket: bad arguments");
  *offset = t->bucket_off[bucket];
  *count = t->bucket_cnt[bucket];
  return GM_OK;
}

const float* gm_train_prediction(const gm_train* t) { return t ? t->pred : nullptr; }

const float* gm_train_activation(const gm_train* t, int32_t which, int32_t step) {
  if (!t) return nullptr;
  const bool step_ok = step >= 0 && step < t->d.m_steps;
  switch (which) {
    case 0: return t->ee_h1;
    case 1: return t->ee_h2;
    case 2: return t->en_h1;
    case 3: return t->en_h2;
    case 4: return step_ok ? t->act[step].eh1 : nullptr;
    case 5: return step_ok ? t->act[step].eh2 : nullptr;
    case 6: return step_ok ? t->act[step].nh1 : nullptr;
    case 7: return step_ok ? t->act[step].nh2 : nullptr;
    case 8: return t->dc_h1;
    case 9: return t->dc_h2;
    default: return nullptr;
  }
}

}  // extern "C"

// =====================================================================================================
// forward + loss + backward
// =====================================================================================================
namespace gm {
namespace {

struct StepCtx {
  gm_train* t;
  cudaStream_t st;
  int N, E;                 // static bounds
  const int* E_dev;         // live edge count
  const int *recv_ptr, *esend, *erecv, *edge_ptr;
  float* P;                 // parameters
  float* G;                 // gradients
};

int gemm(const StepCtx& c, const float* x, const void* chunk, const float* bias, const float* add1, const int* idx1,
         const float* add2, const int* idx2, const float* mask, bool relu, bool accumulate, bool edges, float* out) {
  TrGemmArgs a{x, nullptr, chunk, bias, add1, idx1, add2, idx2, mask, out, relu ? 1 : 0, accumulate ? 1 : 0,
               edges ? c.E : c.N, edges ? c.E_dev : nullptr};
  return tr_gemm(a, c.st);
}
int wgrad(const StepCtx& c, const float* A, const float* B, bool edges, size_t w_off, int ld, int col0, size_t b_off,
          bool with_bias) {
  return tr_wgrad(A, B, edges ? c.E : c.N, edges ? c.E_dev : nullptr, c.G + w_off, ld, col0,
                  with_bias ? c.G + b_off : nullptr, c.t->wg_partial, c.st);
}
#define TR(call) do { int _rc = (call); if (_rc) return _rc; } while (0)

// hidden layers 2, 3 + LayerNorm of one MLP, forward: h1 -> h2 (saved) -> z -> xhat, rstd (saved) -> y / residual
int mlp_tail_fwd(const StepCtx& c, const MlpRef& p, const Block& W2, const Block& W3, bool edges, const float* h1,
                 float* h2, float* z, float* xhat, float* rstd, float* y_out, const float* res_in, float* res_out) {
  TR(gemm(c, h1, W2.fwd, c.P + p.b2, nullptr, nullptr, nullptr, nullptr, nullptr, true, false, edges, h2));
  // layer 3 with the LayerNorm (+ residual) in its epilogue: the pre-normalisation rows z never go to memory
  (void)z;
  TrGemmArgs a{h2, nullptr, W3.fwd, c.P + p.b3, nullptr, nullptr, nullptr, nullptr, nullptr, y_out, 0, 0,
               edges ? c.E : c.N, edges ? c.E_dev : nullptr, c.P + p.g, c.P + p.b, xhat, rstd, res_in, res_out};
  return tr_gemm(a, c.st);
}

// ... and backward: dy (= d1 + d2[idx2]) -> dz3 -> dz2 -> dz1 (returned in dz_a), weight / bias / LayerNorm gradients
int mlp_tail_bwd(const StepCtx& c, const MlpRef& p, const Block& W2, const Block& W3, bool edges, const float* d1,
                 const float* d2, const int* idx2, const float* xhat, const float* rstd, const float* h2,
                 const float* h1, float* dz_a, float* dz_b) {
  const int M = edges ? c.E : c.N;
  const int* Md = edges ? c.E_dev : nullptr;
  TR(tr_ln_bwd(d1, d2, idx2, xhat, rstd, c.P + p.g, M, Md, dz_a, c.G + p.g, c.G + p.b, c.t->ln_partial, c.st));
  TR(wgrad(c, dz_a, h2, edges, p.W3, H, 0, p.b3, true));                                    // dW3, db3
  TR(gemm(c, dz_a, W3.bwd, nullptr, nullptr, nullptr, nullptr, nullptr, h2, false, false, edges, dz_b));   // dz2
  TR(wgrad(c, dz_b, h1, edges, p.W2, H, 0, p.b2, true));                                    // dW2, db2
  TR(gemm(c, dz_b, W2.bwd, nullptr, nullptr, nullptr, nullptr, nullptr, h1, false, false, edges, dz_a));   // dz1
  return GM_OK;
}

int forward(const StepCtx& c, const float* nodes, const float* pos, int pos_stride, float conn_r, const float* target,
            const unsigned char* row_mask, float* loss) {
  gm_train* t = c.t;
  const gm_train_desc& d = t->d;
  // encoders (epd_gnn.py:30-33, 88)
  TR(edge_features(pos, pos_stride, c.esend, c.erecv, c.E_dev, c.E, conn_r, t->ea, c.st));
  TR(tr_small_fwd(t->ea, d.edge_dim, c.P + t->enc_edge.p.W1, c.P + t->enc_edge.p.b1, H, c.E, c.E_dev, 1, t->ee_h1, c.st));
  TR(mlp_tail_fwd(c, t->enc_edge.p, t->enc_edge.W2, t->enc_edge.W3, true, t->ee_h1, t->ee_h2, t->zE, t->ee_xhat,
                  t->ee_rstd, t->act[0].e_in, nullptr, nullptr));
  TR(tr_small_fwd(nodes, d.node_dim, c.P + t->enc_node.p.W1, c.P + t->enc_node.p.b1, H, c.N, nullptr, 1, t->en_h1, c.st));
  TR(mlp_tail_fwd(c, t->enc_node.p, t->enc_node.W2, t->enc_node.W3, false, t->en_h1, t->en_h2, t->zN, t->en_xhat,
                  t->en_rstd, t->act[0].v_in, nullptr, nullptr));
  for (int s = 0; s < d.m_steps; ++s) {
    StepAct& a = t->act[s];
    float* e_next = s + 1 < d.m_steps ? t->act[s + 1].e_in : t->e_last;
    float* v_next = s + 1 < d.m_steps ? t->act[s + 1].v_in : t->v_last;
    const EdgeMlp& pe = t->pe[s];
    const NodeMlp& pn = t->pn[s];
    // edge half (graph_networks.py:43-50): h1 = relu(e We^T + (v Wr^T + b1)[recv] + (v Ws^T)[send])
    TR(gemm(c, a.v_in, pe.Wr.fwd, c.P + pe.p.b1, nullptr, nullptr, nullptr, nullptr, nullptr, false, false, false, t->Pr));
    TR(gemm(c, a.v_in, pe.Ws.fwd, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, false, false, false, t->Ps));
    TR(gemm(c, a.e_in, pe.We.fwd, nullptr, t->Pr, c.erecv, t->Ps, c.esend, nullptr, true, false, true, a.eh1));
    TR(mlp_tail_fwd(c, pe.p, pe.W2, pe.W3, true, a.eh1, a.eh2, t->zE, a.exhat, a.erstd, t->yE, a.e_in, e_next));
    // node half (graph_networks.py:75-85) on the PRE-residual edge output; residuals epd_gnn.py:103-104
    TR(segment_sum(t->yE, H, c.recv_ptr, nullptr, c.N, a.agg, c.st));
    TR(gemm(c, a.v_in, pn.W1b.fwd, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, false, false, false, t->tN));
    TR(gemm(c, a.agg, pn.W1a.fwd, c.P + pn.p.b1, t->tN, nullptr, nullptr, nullptr, nullptr, true, false, false, a.nh1));
    TR(mlp_tail_fwd(c, pn.p, pn.W2, pn.W3, false, a.nh1, a.nh2, t->zN, a.nxhat, a.nrstd, nullptr, a.v_in, v_next));
  }
  // decoder (epd_gnn.py:49, 96)
  const DecMlp& dc = t->dec;
  TR(gemm(c, t->v_last, dc.W1.fwd, c.P + dc.p.b1, nullptr, nullptr, nullptr, nullptr, nullptr, true, false, false, t->dc_h1));
  TR(gemm(c, t->dc_h1, dc.W2.fwd, c.P + dc.p.b2, nullptr, nullptr, nullptr, nullptr, nullptr, true, false, false, t->dc_h2));
  TR(tr_small_fwd(t->dc_h2, H, c.P + dc.p.W3, c.P + dc.p.b3, d.out_dim, c.N, nullptr, 0, t->pred, c.st));
  if (target != nullptr)
    TR(tr_l1_loss(t->pred, target, row_mask, c.N, d.out_dim, loss, t->dpred, t->loss_scratch, c.st));
  return GM_OK;
}

int backward(const StepCtx& c, const float* nodes, const int* eref, gm_grad_ready_fn cb, void* user) {
  gm_train* t = c.t;
  const gm_train_desc& d = t->d;
  GM_CUDA(cudaMemsetAsync(c.G, 0, sizeof(float) * t->n_params, c.st));
  GM_CUDA(cudaMemsetAsync(t->de, 0, sizeof(float) * (size_t)c.E * H, c.st));   // nothing consumes e_m: its gradient is zero
  invert_perm_kernel<<<ceil_div(c.E, 256), 256, 0, c.st>>>(eref, c.E, c.E_dev, t->inv_eref);
  GM_LAUNCHED(1);
  int bucket = 0;
  auto ready = [&]() { if (cb) cb(user, bucket, t->bucket_off[bucket], t->bucket_cnt[bucket]); ++bucket; };
  // decoder: pred = h2 W3^T + b3 (CUDA cores), h2 = relu(.), h1 = relu(.)
  const DecMlp& dc = t->dec;
  TR(tr_small_bwd(t->dc_h2, H, c.P + dc.p.W3, d.out_dim, t->dpred, nullptr, c.N, nullptr, c.G + dc.p.W3, c.G + dc.p.b3,
                  t->dzN_a, t->dc_h2, t->small_partial, c.st));                                // dz2 = (dpred W3) o (h2 > 0)
  TR(wgrad(c, t->dzN_a, t->dc_h1, false, dc.p.W2, H, 0, dc.p.b2, true));
  TR(gemm(c, t->dzN_a, dc.W2.bwd, nullptr, nullptr, nullptr, nullptr, nullptr, t->dc_h1, false, false, false, t->dzN_b));  // dz1
  TR(wgrad(c, t->dzN_b, t->v_last, false, dc.p.W1, H, 0, dc.p.b1, true));
  TR(gemm(c, t->dzN_b, dc.W1.bwd, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, false, false, false, t->dv));      // d v_m
  ready();
  for (int s = d.m_steps - 1; s >= 0; --s) {
    StepAct& a = t->act[s];
    const EdgeMlp& pe = t->pe[s];
    const NodeMlp& pn = t->pn[s];
    // node half: v_{s+1} = v_s + LN(.)  ->  dy = dv
    TR(mlp_tail_bwd(c, pn.p, pn.W2, pn.W3, false, t->dv, nullptr, nullptr, a.nxhat, a.nrstd, a.nh2, a.nh1, t->dzN_a, t->dzN_b));
    TR(wgrad(c, t->dzN_a, a.agg, false, pn.p.W1, 2 * H, 0, pn.p.b1, true));                     // dW1[:, :128], db1
    TR(wgrad(c, t->dzN_a, a.v_in, false, pn.p.W1, 2 * H, H, 0, false));                         // dW1[:, 128:]
    TR(gemm(c, t->dzN_a, pn.W1a.bwd, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, false, false, false, t->dagg));
    TR(gemm(c, t->dzN_a, pn.W1b.bwd, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, false, true, false, t->dv));
    // edge half: e_{s+1} = e_s + y, agg = segment_sum(y)  ->  dy = de + dagg[recv]
    TR(mlp_tail_bwd(c, pe.p, pe.W2, pe.W3, true, t->de, t->dagg, c.erecv, a.exhat, a.erstd, a.eh2, a.eh1, t->dzE_a, t->dzE_b));
    TR(wgrad(c, t->dzE_a, a.e_in, true, pe.p.W1, 3 * H, 0, pe.p.b1, true));                     // dWe, db1 (b1 rides on Pr[recv])
    TR(gemm(c, t->dzE_a, pe.We.bwd, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, false, true, true, t->de));
    TR(segment_sum(t->dzE_a, H, c.recv_ptr, nullptr, c.N, t->dPr, c.st));                       // gather^T over receivers
    TR(segment_sum(t->dzE_a, H, c.edge_ptr, t->inv_eref, c.N, t->dPs, c.st));                   // gather^T over senders
    TR(wgrad(c, t->dPr, a.v_in, false, pe.p.W1, 3 * H, H, 0, false));                           // dWr
    TR(wgrad(c, t->dPs, a.v_in, false, pe.p.W1, 3 * H, 2 * H, 0, false));                       // dWs
    TR(gemm(c, t->dPr, pe.Wr.bwd, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, false, true, false, t->dv));
    TR(gemm(c, t->dPs, pe.Ws.bwd, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, false, true, false, t->dv));
    ready();
  }
  // encoders: e_0 = LN(.), v_0 = LN(.)
  TR(mlp_tail_bwd(c, t->enc_edge.p, t->enc_edge.W2, t->enc_edge.W3, true, t->de, nullptr, nullptr, t->ee_xhat,
                  t->ee_rstd, t->ee_h2, t->ee_h1, t->dzE_a, t->dzE_b));
  TR(tr_small_bwd(t->ea, d.edge_dim, c.P + t->enc_edge.p.W1, H, t->dzE_a, nullptr, c.E, c.E_dev, c.G + t->enc_edge.p.W1,
                  c.G + t->enc_edge.p.b1, nullptr, nullptr, t->small_partial, c.st));
  TR(mlp_tail_bwd(c, t->enc_node.p, t->enc_node.W2, t->enc_node.W3, false, t->dv, nullptr, nullptr, t->en_xhat,
                  t->en_rstd, t->en_h2, t->en_h1, t->dzN_a, t->dzN_b));
  TR(tr_small_bwd(nodes, d.node_dim, c.P + t->enc_node.p.W1, H, t->dzN_a, nullptr, c.N, nullptr, c.G + t->enc_node.p.W1,
                  c.G + t->enc_node.p.b1, nullptr, nullptr, t->small_partial, c.st));
  ready();
  return GM_OK;
}

}  // namespace
}  // namespace gm

extern "C" {

int gm_train_forward_backward(gm_train* t, const float* nodes, const float* pos, int32_t pos_stride, double conn_r,
                              int32_t n_nodes, int32_t max_edges, const int32_t* n_edges_dev, const int32_t* recv_ptr,
                              const int32_t* esend, const int32_t* erecv, const int32_t* eref, const int32_t* edge_ptr,
                              const float* target, const uint8_t* row_mask, float* loss, int32_t with_backward,
                              gm_grad_ready_fn grad_ready, void* user, void* stream) {
  GM_REQUIRE(t != nullptr && nodes != nullptr && pos != nullptr && n_edges_dev != nullptr && recv_ptr != nullptr &&
                 esend != nullptr && erecv != nullptr, "gm_train_forward_backward: NULL argument");
  GM_REQUIRE(n_nodes > 0 && n_nodes <= t->d.max_nodes && max_edges > 0 && max_edges <= t->d.max_edges,
             "gm_train_forward_backward: batch of %d nodes / %d edge slots exceeds the plan (%d / %d)", n_nodes,
             max_edges, t->d.max_nodes, t->d.max_edges);
  GM_REQUIRE(t->d.edge_dim == 4, "the training step builds 4-wide displacement edge features (edge_dim=%d)", t->d.edge_dim);
  GM_REQUIRE(!with_backward || (target != nullptr && loss != nullptr && eref != nullptr && edge_ptr != nullptr),
             "gm_train_forward_backward: the backward needs target, loss, eref and edge_ptr");
  StepCtx c{t, (cudaStream_t)stream, n_nodes, max_edges, n_edges_dev, recv_ptr, esend, erecv, edge_ptr, t->d.params,
            t->d.grads};
  int rc = forward(c, nodes, pos, pos_stride, (float)conn_r, target, row_mask, loss);
  if (rc || !with_backward) return rc;
  return backward(c, nodes, eref, grad_ready, user);
}

}  // extern "C"
