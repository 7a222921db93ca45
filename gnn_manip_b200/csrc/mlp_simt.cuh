#pragma once
#include "common.cuh"

namespace gm {

constexpr int kMaxSimtLayers = 8;

struct MlpSimtLayer {
  const float* Wt;  // [in][out] fp32 (transposed nn.Linear weight)
  const float* b;   // [out] or null
  int in, out;
};

struct MlpSimtArgs {
  int n_layers;
  MlpSimtLayer layers[kMaxSimtLayers];   // ReLU between layers, none after the last
  const float* ln_g;                     // LayerNorm affine (null = no norm)
  const float* ln_b;
  float ln_eps;
  int n_src;                             // input row m = concat_s src[s][idx[s] ? idx[s][m] : m]
  const float* src[3];
  const int* idx[3];
  int width[3];
  int src_stride[3];
  int M;                                 // rows (upper bound)
  const int* M_dev;                      // optional live row count on device
  const int* out_perm;                   // optional: row m is written to out row out_perm[m]
  const float* residual;                 // optional, same indexing as out
  float* out;                            // y (+ residual)
  float* out_pre;                        // optional y before the residual
  int max_width;                         // filled by the launcher
};

int mlp_simt_configure();
int mlp_simt(const MlpSimtArgs& args, cudaStream_t stream);
int segment_sum(const float* src, int F, const int* recv_ptr, const int* eref, int n_nodes, float* agg,
                cudaStream_t stream);
int gather_rows(const float* src, int F, const int* idx, long long M, float* out, cudaStream_t stream);

}  // namespace gm
