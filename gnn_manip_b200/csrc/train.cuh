// Rows E(2) + N3 -- building blocks of the training step (forward with saved activations, backward, loss, Adam).
#pragma once
#include "common.cuh"

namespace gm {

constexpr int kTrH = 128;                   // hidden width served by the tensor-core training kernels
constexpr int kTrChunkBytes = 65536;        // packed 128 x 128 operand: bf16 hi image + bf16 lo image (K-major, SW128)

// out[r, :] = epilogue(X[in_idx ? in_idx[r] : r, :] @ B^T) for r < M (M optionally read from the device),
//   B: packed chunk of a [128, 128] matrix (tr_pack: rows = output features, K-major), bf16x3 on tcgen05
//   v = acc + bias + add1[idx1 ? idx1[r] : r] + add2[idx2 ? idx2[r] : r]      (each term optional)
//   v = relu ? max(v, 0) : v
//   v = mask ? (mask[r] > 0 ? v : 0) : v                                        (ReLU backward)
//   out[r] = accumulate ? out[r] + v : v
struct TrGemmArgs {
  const float* x;
  const int* in_idx;
  const void* chunk;
  const float* bias;
  const float* add1;
  const int* idx1;
  const float* add2;
  const int* idx2;
  const float* mask;
  float* out;
  int relu;
  int accumulate;
  int M;
  const int* M_dev;
  // LayerNorm epilogue (ln_g != null; excludes add1 / add2 / relu / mask / accumulate): z = acc + bias is normalised
  // per row, xhat / rstd saved (optional), out (optional) = xhat * g + b, res_out (optional) = res_in + out
  const float* ln_g;
  const float* ln_b;
  float* ln_xhat;
  float* ln_rstd;
  const float* res_in;
  float* res_out;
};
int tr_gemm(const TrGemmArgs& a, cudaStream_t stream);

// dW[128, ld] columns [col0, col0 + 128) += A^T B over rows r < M (A, B: [M, 128] fp32), optionally
// dbias[128] += column sums of A.  Deterministic: per-CTA partials in `partial` reduced in a fixed order.
// partial: workspace of tr_wgrad_partial_floats() floats.
size_t tr_wgrad_partial_floats();
int tr_wgrad(const float* A, const float* B, int M, const int* M_dev, float* dW, int ld, int col0, float* dbias,
             float* partial, cudaStream_t stream);

// w [128 rows][ld] columns [col0, col0+128) -> chunk (transpose = 0: B = w, for y = x w^T; 1: B = w^T, for dx = dy w)
struct TrPackJob {
  const float* w;
  int ld, col0, transpose;
  void* chunk;
};
int tr_pack(const TrPackJob* jobs_dev, int n_jobs, cudaStream_t stream);

// LayerNorm forward over rows of 128: xhat = (z - mean) rstd, y = xhat g + b; y_out / res_out = res_in + y optional
int tr_ln_fwd(const float* z, const float* g, const float* b, int M, const int* M_dev, float* xhat, float* rstd,
              float* y_out, const float* res_in, float* res_out, cudaStream_t stream);
// LayerNorm backward: dy[r] = d1[r] + (d2 ? d2[idx2 ? idx2[r] : r] : 0); dz = rstd (g dy - mean(g dy) - xhat mean(g dy xhat));
// dgamma += sum dy xhat, dbeta += sum dy (deterministic two-stage; partial: tr_ln_partial_floats(M) floats)
size_t tr_ln_partial_floats(int M);
int tr_ln_bwd(const float* d1, const float* d2, const int* idx2, const float* xhat, const float* rstd, const float* g,
              int M, const int* M_dev, float* dz, float* dgamma, float* dbeta, float* partial, cudaStream_t stream);
// the same with the partials in a file-scope device array (no scratch argument, at most 65 536 rows; train_kernels.cu)
int tr_ln_bwd_static(const float* d1, const float* d2, const int* idx2, const float* xhat, const float* rstd,
                     const float* g, int M, float* dz, float* dgamma, float* dbeta, cudaStream_t stream);

// Small first / last layers on CUDA cores.
//   fwd:  h[r, o] = act(x[r, :K] . w[o, :K] + b[o])   (w [O][K] row-major, K, O <= 128)
//   bwd:  dz = mask ? dh (mask > 0) : dh;  dW[o, k] += sum_r dz[r, o] x[r, k];  db[o] += sum_r dz[r, o];
//         dx[r, k] = sum_o dz[r, o] w[o, k] (optional), zeroed where dx_mask[r, k] <= 0 (optional: ReLU of the layer below)
int tr_small_fwd(const float* x, int K, const float* w, const float* b, int O, int M, const int* M_dev, int relu,
                 float* h, cudaStream_t stream);
size_t tr_small_partial_floats(int M, int K, int O);
int tr_small_bwd(const float* x, int K, const float* w, int O, const float* dh, const float* mask, int M,
                 const int* M_dev, float* dW, float* db, float* dx, const float* dx_mask, float* partial,
                 cudaStream_t stream);

// L1Loss(reduction='sum') / n_rows (examples/train_dyn.py:56-64, 109): loss scalar (device) and dpred = sign(pred - tgt) / n
// over the rows selected by row_mask (null = all rows; the `use_updated_loss` variant keeps sand rows only)
int tr_l1_loss(const float* pred, const float* tgt, const unsigned char* row_mask, int M, int width, float* loss,
               float* dpred, float* scratch /*>= 2 + ceil(M/256) floats*/, cudaStream_t stream);

// torch.optim.Adam (lr, betas, eps; no weight decay, no amsgrad) on a flat parameter vector; step = 1, 2, ...
int tr_adam(float* p, const float* g, float* m, float* v, size_t n, float lr, float beta1, float beta2, float eps,
            int step, cudaStream_t stream);

// out[r] = a[r] + b[idx ? idx[r] : r] (rows of 128) -- gradient joins
int tr_add_rows(const float* a, const float* b, const int* idx, int M, const int* M_dev, float* out, cudaStream_t stream);

}  // namespace gm
