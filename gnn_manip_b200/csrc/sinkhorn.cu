// Row N1 -- batched Sinkhorn divergence between the rolled-out sand clouds of a CMA-ES population and the target
// cloud: the objective tail of CMAESolver.cma_objective (gnn_manip/utils/traj_utils.py:154-158), where the reference
// calls geomloss.SamplesLoss(loss="sinkhorn", p=2, blur=.05) once per candidate on the CPU (traj_utils.py:69,
// 276, 439).
//
// geomloss is not part of the reference checkout (unpinned pip dependency, environment.yml:25) -> the algorithm is
// written after its published tensorized implementation (sinkhorn_divergence.py: epsilon_schedule, sinkhorn_loop with
// symmetrised updates, debiased sinkhorn_cost; sinkhorn_samples.py: softmin_tensorized); the tests check it against
// an fp64 numpy statement of the same algorithm (oracle/).  PARITY UNPINNED against geomloss itself.
//
// One candidate = four coupled log-sum-exp problems per iteration (f_ba over the target cloud, g_ab over the source
// cloud, and the two debiasing self terms); the cost matrix |x - y|^2 / 2 is never stored: a thread owns one row,
// the column cloud streams through shared memory in tiles, and the row's log-sum-exp is accumulated online in base 2.
#include "sinkhorn.cuh"

namespace gm {

namespace {

constexpr int kSinkThreads = 256;
constexpr int kSinkTile = 1024;              // columns per shared-memory tile (16 KB)
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;

__global__ void sink_pack_kernel(const float* __restrict__ x, int x_stride, long long x_batch_stride,
                                 const int* __restrict__ x_idx, int n, const float* __restrict__ y, int m, int n_batch,
                                 float4* __restrict__ xs, float4* __restrict__ ys) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long nx = (long long)n_batch * n;
  if (i < nx) {
    const int b = (int)(i / n), k = (int)(i - (long long)b * n);
    const float* p = x + (size_t)b * x_batch_stride + (size_t)(x_idx ? x_idx[k] : k) * x_stride;
    xs[i] = make_float4(p[0], p[1], p[2], 0.f);
  } else if (i < nx + m) {
    const float* p = y + (size_t)(i - nx) * 3;
    ys[i - nx] = make_float4(p[0], p[1], p[2], 0.f);
  }
}

// geomloss scaling_parameters / max_diameter / epsilon_schedule (sinkhorn_divergence.py), one block per candidate:
//   diameter = ||max(x, y) - min(x, y)||_2 over the joint bounding box (fp32, as torch computes it),
//   eps_list = [diameter^2] + [exp(e) for e in arange(2 log diameter, 2 log blur, 2 log scaling)] + [blur^2]
__global__ void __launch_bounds__(kSinkThreads)
sink_schedule_kernel(const float4* __restrict__ xs, int n, const float4* __restrict__ ys, int m, double blur,
                     double scaling, float* __restrict__ eps, int* __restrict__ n_eps) {
  __shared__ float s_mn[3][kSinkThreads / 32], s_mx[3][kSinkThreads / 32];
  const int b = blockIdx.x;
  float mn[3] = {3.4e38f, 3.4e38f, 3.4e38f}, mx[3] = {-3.4e38f, -3.4e38f, -3.4e38f};
  for (int i = threadIdx.x; i < n + m; i += blockDim.x) {
    const float4 p = i < n ? xs[(size_t)b * n + i] : ys[i - n];
    mn[0] = fminf(mn[0], p.x); mn[1] = fminf(mn[1], p.y); mn[2] = fminf(mn[2], p.z);
    mx[0] = fmaxf(mx[0], p.x); mx[1] = fmaxf(mx[1], p.y); mx[2] = fmaxf(mx[2], p.z);
  }
#pragma unroll
  for (int d = 0; d < 3; ++d) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      mn[d] = fminf(mn[d], __shfl_xor_sync(0xffffffffu, mn[d], o));
      mx[d] = fmaxf(mx[d], __shfl_xor_sync(0xffffffffu, mx[d], o));
    }
    if ((threadIdx.x & 31) == 0) { s_mn[d][threadIdx.x >> 5] = mn[d]; s_mx[d][threadIdx.x >> 5] = mx[d]; }
  }
  __syncthreads();
  if (threadIdx.x != 0) return;
  float ext2 = 0.f;
  for (int d = 0; d < 3; ++d) {
    float lo = s_mn[d][0], hi = s_mx[d][0];
    for (int w = 1; w < kSinkThreads / 32; ++w) { lo = fminf(lo, s_mn[d][w]); hi = fmaxf(hi, s_mx[d][w]); }
    const float e = hi - lo;
    ext2 += e * e;
  }
  const double diameter = (double)sqrtf(ext2);
  float* out = eps + (size_t)b * kSinkMaxEps;
  int k = 0;
  out[k++] = (float)(diameter * diameter);
  const double start = 2.0 * log(diameter), stop = 2.0 * log(blur), step = 2.0 * log(scaling);
  if (diameter > 0.0 && step < 0.0) {
    const int cnt = (int)ceil((stop - start) / step);          // numpy.arange length
    for (int j = 0; j < cnt && k < kSinkMaxEps - 1; ++j) out[k++] = (float)exp(start + j * step);
  }
  out[k++] = (float)(blur * blur);
  n_eps[b] = k;
}

// mode 0: initialisation at eps_list[0] (no dual term, no averaging); mode 1: iteration `it` (symmetrised update);
// mode 2: final extrapolation at eps_list[-1] (no averaging).  Problems (blockIdx.y): 0 f_ba rows x / cols y with
// g_ab; 1 g_ab rows y / cols x with f_ba; 2 f_aa rows x / cols x with f_aa; 3 g_bb rows y / cols y with g_bb.
__global__ void __launch_bounds__(kSinkThreads)
sink_softmin_kernel(const float4* __restrict__ xs, int n, const float4* __restrict__ ys, int m, int P, int n_batch,
                    const float* __restrict__ eps_all, const int* __restrict__ n_eps, int mode, int it,
                    const float* __restrict__ pot_in, float* __restrict__ pot_out) {
  __shared__ float4 s_col[kSinkTile];
  const int q = blockIdx.y, b = blockIdx.z;
  const bool rows_x = (q == 0 || q == 2), cols_x = (q == 1 || q == 2);
  const int nr = rows_x ? n : m, nc = cols_x ? n : m;
  const int row0 = blockIdx.x * kSinkThreads;
  if (row0 >= nr) return;
  const int src_pot = q == 0 ? 1 : (q == 1 ? 0 : q);          // which potential rides on the columns
  const size_t plane = (size_t)n_batch * P;
  const float* pin_cols = pot_in + (size_t)src_pot * plane + (size_t)b * P;
  const float* pin_rows = pot_in + (size_t)q * plane + (size_t)b * P;
  float* pout = pot_out + (size_t)q * plane + (size_t)b * P;
  const int row = row0 + threadIdx.x;
  const int ne = n_eps[b];
  if (mode == 1 && it >= ne) {                                 // this candidate's schedule is shorter: carry over
    if (row < nr) pout[row] = pin_rows[row];
    return;
  }
  const float eps = eps_all[(size_t)b * kSinkMaxEps + (mode == 0 ? 0 : (mode == 1 ? it : ne - 1))];
  const float inv_eps = 1.0f / eps;
  const float logw = -logf((float)nc);                         // uniform weights 1 / nc
  const float kcoef = -0.5f * inv_eps * kLog2e;                // -(|x - y|^2 / 2) / eps in base 2
  const float4* cols = cols_x ? xs + (size_t)b * n : ys;
  float4 me = make_float4(0.f, 0.f, 0.f, 0.f);
  if (row < nr) me = rows_x ? xs[(size_t)b * n + row] : ys[row];
  float run_m = -3.0e38f, run_s = 0.f;
  for (int c0 = 0; c0 < nc; c0 += kSinkTile) {
    const int tn = min(kSinkTile, nc - c0);
    __syncthreads();
    for (int j = threadIdx.x; j < tn; j += kSinkThreads) {
      float4 p = cols[c0 + j];
      const float h = mode == 0 ? logw : logw + pin_cols[c0 + j] * inv_eps;
      p.w = h * kLog2e;
      s_col[j] = p;
    }
    __syncthreads();
    float tm = -3.0e38f;
#pragma unroll 4
    for (int j = 0; j < tn; ++j) {
      const float4 p = s_col[j];
      const float dx = me.x - p.x, dy = me.y - p.y, dz = me.z - p.z;
      tm = fmaxf(tm, fmaf(kcoef, fmaf(dz, dz, fmaf(dy, dy, dx * dx)), p.w));
    }
    const float nm = fmaxf(run_m, tm);
    float ts = 0.f;
#pragma unroll 4
    for (int j = 0; j < tn; ++j) {
      const float4 p = s_col[j];
      const float dx = me.x - p.x, dy = me.y - p.y, dz = me.z - p.z;
      ts += exp2f(fmaf(kcoef, fmaf(dz, dz, fmaf(dy, dy, dx * dx)), p.w) - nm);
    }
    run_s = run_s * exp2f(run_m - nm) + ts;
    run_m = nm;
  }
  if (row < nr) {
    const float out = -eps * kLn2 * (run_m + log2f(run_s));
    pout[row] = mode == 1 ? 0.5f * (pin_rows[row] + out) : out;
  }
}

// sinkhorn_cost, debiased, no unbalanced term: sum_i a_i (f_ba - f_aa)_i + sum_j b_j (g_ab - g_bb)_j
__global__ void __launch_bounds__(kSinkThreads)
sink_cost_kernel(const float* __restrict__ pot, int n, int m, int P, int n_batch, float* __restrict__ loss) {
  __shared__ float s_red[kSinkThreads / 32];
  const int b = blockIdx.x;
  const size_t plane = (size_t)n_batch * P;
  const float* f_ba = pot + (size_t)b * P;
  const float* g_ab = pot + plane + (size_t)b * P;
  const float* f_aa = pot + 2 * plane + (size_t)b * P;
  const float* g_bb = pot + 3 * plane + (size_t)b * P;
  float acc = 0.f;
  for (int i = threadIdx.x; i < n; i += blockDim.x) acc += (f_ba[i] - f_aa[i]) / (float)n;
  for (int j = threadIdx.x; j < m; j += blockDim.x) acc += (g_ab[j] - g_bb[j]) / (float)m;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int w = 0; w < kSinkThreads / 32; ++w) t += s_red[w];
    loss[b] = t;
  }
}

}  // namespace

void sinkhorn_carve(Arena& a, int n_batch, int n, int m, SinkhornWs& ws) {
  const size_t P = (size_t)(n > m ? n : m);
  ws.xs = a.take<float>((size_t)n_batch * n * 4);
  ws.ys = a.take<float>((size_t)m * 4);
  ws.pot = a.take<float>(2 * 4 * (size_t)n_batch * P);
  ws.eps = a.take<float>((size_t)n_batch * kSinkMaxEps);
  ws.n_eps = a.take<int>(n_batch);
}

size_t sinkhorn_workspace_bytes(int n_batch, int n, int m) {
  Arena a(nullptr, 0);
  SinkhornWs ws;
  sinkhorn_carve(a, n_batch, n, m, ws);
  return a.off + 256;
}

int sinkhorn_divergence(const float* x, int x_stride, long long x_batch_stride, const int* x_idx, int n,
                        const float* y, int m, int n_batch, double blur, double scaling, int max_iters, float* loss,
                        const SinkhornWs& ws, cudaStream_t stream) {
  GM_REQUIRE(n > 0 && m > 0 && n_batch > 0, "sinkhorn_divergence: empty clouds (n=%d m=%d batch=%d)", n, m, n_batch);
  GM_REQUIRE(blur > 0.0 && scaling > 0.0 && scaling < 1.0, "sinkhorn_divergence: blur=%g scaling=%g", blur, scaling);
  GM_REQUIRE(max_iters >= 2 && max_iters <= kSinkMaxEps, "sinkhorn_divergence: max_iters=%d outside [2,%d]", max_iters,
             kSinkMaxEps);
  const int P = n > m ? n : m;
  float4* xs = reinterpret_cast<float4*>(ws.xs);
  float4* ys = reinterpret_cast<float4*>(ws.ys);
  const long long total = (long long)n_batch * n + m;
  sink_pack_kernel<<<ceil_div(total, 256), 256, 0, stream>>>(x, x_stride, x_batch_stride, x_idx, n, y, m, n_batch, xs, ys);
  sink_schedule_kernel<<<n_batch, kSinkThreads, 0, stream>>>(xs, n, ys, m, blur, scaling, ws.eps, ws.n_eps);
  const size_t buf = 4 * (size_t)n_batch * P;
  const dim3 grid(ceil_div(P, kSinkThreads), 4, n_batch);
  int cur = 0;
  sink_softmin_kernel<<<grid, kSinkThreads, 0, stream>>>(xs, n, ys, m, P, n_batch, ws.eps, ws.n_eps, 0, 0,
                                                       ws.pot + (size_t)(cur ^ 1) * buf, ws.pot + (size_t)cur * buf);
  for (int it = 0; it < max_iters; ++it) {
    sink_softmin_kernel<<<grid, kSinkThreads, 0, stream>>>(xs, n, ys, m, P, n_batch, ws.eps, ws.n_eps, 1, it,
                                                         ws.pot + (size_t)cur * buf, ws.pot + (size_t)(cur ^ 1) * buf);
    cur ^= 1;
  }
  sink_softmin_kernel<<<grid, kSinkThreads, 0, stream>>>(xs, n, ys, m, P, n_batch, ws.eps, ws.n_eps, 2, 0,
                                                       ws.pot + (size_t)cur * buf, ws.pot + (size_t)(cur ^ 1) * buf);
  cur ^= 1;
  sink_cost_kernel<<<n_batch, kSinkThreads, 0, stream>>>(ws.pot + (size_t)cur * buf, n, m, P, n_batch, loss);
  GM_LAUNCHED(5 + max_iters);
  GM_LAUNCH_CHECK("sinkhorn kernels");
  return GM_OK;
}

}  // namespace gm
