#pragma once
#include "common.cuh"

namespace gm {

constexpr int kSinkMaxEps = 24;   // entries of the epsilon-scaling schedule kept per candidate

struct SinkhornWs {
  float* xs;        // [B, N, 4] packed source points (x, y, z, 0)
  float* ys;        // [M, 4] packed target points
  float* pot;       // [2][4][B][P] potentials (f_ba, g_ab, f_aa, g_bb), double buffered; P = max(N, M)
  float* eps;       // [B, kSinkMaxEps] schedule (eps values, fp32 like geomloss' python floats cast in torch ops)
  int* n_eps;       // [B]
};

size_t sinkhorn_workspace_bytes(int n_batch, int n, int m);
void sinkhorn_carve(Arena& a, int n_batch, int n, int m, SinkhornWs& ws);

// Debiased Sinkhorn divergence S_eps(alpha, beta) between uniform point clouds, p = 2, cost |x - y|^2 / 2,
// epsilon scaling from diameter^2 down to blur^2 with ratio scaling^2, symmetrised updates -- the algorithm of
// geomloss.SamplesLoss("sinkhorn", p=2, blur, scaling) in its tensorized form.
//   x: source points of candidate b, point i at x + b * x_batch_stride + (x_idx ? x_idx[i] : i) * x_stride (3 floats)
//   y: [m, 3] target cloud shared by all candidates
//   max_iters: upper bound on the schedule length (from the largest possible diameter); loss [n_batch]
int sinkhorn_divergence(const float* x, int x_stride, long long x_batch_stride, const int* x_idx, int n,
                        const float* y, int m, int n_batch, double blur, double scaling, int max_iters, float* loss,
                        const SinkhornWs& ws, cudaStream_t stream);

}  // namespace gm
