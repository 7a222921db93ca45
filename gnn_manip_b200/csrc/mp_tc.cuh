#pragma once
#include "common.cuh"
#include "physics.cuh"

namespace gm {

constexpr int kH = 128;                  // hidden size served by the tcgen05 path
constexpr int kTile = 128;               // rows per UMMA tile (M = 128, cta_group::1)
constexpr int kAggRows = 64;             // rows per aggregation sub-tile (carry granularity)
constexpr int kChunkBytes = 65536;       // one packed 128x128 weight chunk: bf16 hi image + bf16 lo image
constexpr int kImageBytes = 32768;       // one 128 x 128 bf16 operand image (two 64-wide K blocks)
constexpr int kKBlockBytes = 16384;      // 128 rows x 128 B (SWIZZLE_128B, K-major)

// Parameters of one 3-layer MLP whose hidden layers run on the tensor cores.
struct TcMlp {
  const float* w1;     // first layer, fp32 [128][in1] row-major (SIMT first layers: encoders)  or null
  const float* b1;     // [128]
  // packed 64 KB chunks (device).  edge MLP: {We, W2, W3, Wr, Ws} (W1 = [We | Wr | Ws]);
  // node MLP: {W1[:, :128] (aggregate), W1[:, 128:] (node), W2, W3}; encoders: {-, -, W2, W3};
  // decoder: {W1, W2}.
  const void* wc[5];
  const float* b2;     // [128]
  const float* b3;     // [128] (decoder: [out_dim])
  const float* ln_g;   // [128] or null
  const float* ln_b;
  const float* w3;     // decoder last layer fp32 [out_dim][128] or null
};

struct TcGraph {
  const int* n_edges_dev;   // live edge count (device)
  int max_edges;
  int n_nodes;
  const int* esend;         // [E] receiver-major order
  const int* erecv;
  const int* recv_ptr;      // [n_nodes + 1]
};

struct TcBuffers {
  float* e;        // [max_edges, 128] edge latents, receiver-major order, updated in place
  float* v;        // [n_nodes, 128] node latents, updated in place
  float* pr;       // [n_nodes, 128] v @ W1[:, 128:256]^T + b1 of the *next* edge MLP (receiver term)
  float* ps;       // [n_nodes, 128] v @ W1[:, 256:384]^T (sender term)
  float* aggpart;  // [n_nodes, 128] per-receiver partial sums (segment part inside its first agg tile)
  float* carry;    // [ceil(max_edges/64), 128] leading-rows partial of every 64-row agg tile
};

int tc_configure();
int tc_debug_stamps(unsigned long long* out_host, int n);   // -DGM_STAMPS builds only
int tc_debug_stamps_edge(unsigned long long* out_host, int n);
int tc_debug_stamps_node16(unsigned long long* out_host, int n);
int tc_enc_edge(const TcMlp& m, int edge_dim, const float* pos, int pos_stride, float conn_r,
                const float* edge_attr /*optional explicit [E,edge_dim] in receiver-major order*/,
                const int* eref /*optional: row map into edge_attr*/, const TcGraph& g, const TcBuffers& b,
                cudaStream_t stream);
int tc_enc_edge_pair(const TcMlp& m, const float* pos, int pos_stride, float conn_r, const float* edge_attr,
                     const int* eref, const TcGraph& g, const TcBuffers& b, cudaStream_t stream);
int tc_enc_node(const TcMlp& m, const TcMlp& next_edge, int node_dim, const float* nodes, const TcGraph& g,
                const TcBuffers& b, cudaStream_t stream);
int tc_mp_edge(const TcMlp& m, const TcGraph& g, const TcBuffers& b, cudaStream_t stream, bool programmatic = true);
// next_edge: the following step's edge MLP (for pr/ps) or, when dec != null, the decoder.
int tc_mp_node(const TcMlp& m, const TcMlp* next_edge, const TcMlp* dec, int out_dim, float* pred,
               const TcGraph& g, const TcBuffers& b, cudaStream_t stream);

// The same node half with 16 warps per tile (thread = row x 32 columns; mp_node16.cu): shorter epilogues per GEMM.
int tc_configure_node16();
int tc_mp_node16(const TcMlp& m, const TcMlp* next_edge, const TcMlp* dec, int out_dim, float* pred, const TcGraph& g,
                 const TcBuffers& b, cudaStream_t stream);

// Pack W[n_out=128][ld] columns [col0, col0+128) (fp32, row-major) into one 64 KB chunk image.
int tc_pack_chunk(const float* w, int ld, int col0, void* chunk, cudaStream_t stream);

// Plain GEMM self-test: out[M,128] = x[M,128] @ W^T (bf16x3 split), M multiple of 128.
int tc_selftest_gemm(const float* x, const void* chunk, float* out, int M, cudaStream_t stream);
// Same product through the tensor-memory operand path of mp_edge.cu; accumulates into out (zero it first).
int tc_selftest_gemm_ts(const float* x, const void* chunk, float* out, int M, cudaStream_t stream);

}  // namespace gm
