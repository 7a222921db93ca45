// Row N2 -- CMA-ES trajectory parametrisations on the device, one block per candidate.
//
// Replaces TrajectoryCMAsolver.interpolate_trajectory (gnn_manip/utils/traj_utils.py:206-228: scale, clip and
// sequential accumulation of per-step increments) and InterpolatedCMAsolver.interpolate_trajectory
// (traj_utils.py:393-416: knots every nr_traj_points steps, scipy.interpolate.pchip_interpolate), plus the curve
// regularisers of compute_loss (compute_vel_acc :161-165, compute_vel_loss / compute_acc_loss :230-246 and :353-373,
// compute_boundaries_penalty :263-274, compute_vel_noninterp :418-436).  The curves land straight in the rot / ty
// buffers of a rollout plan (physics.cu turns them into rigid-particle targets, traj_utils.py:167-194).
//
// All arithmetic is fp64 in numpy's / scipy's operation order with no FMA contraction (__dmul_rn / __dadd_rn ...):
// the curves are bit-identical to the reference's (tests/golden/traj_param.npz); the norm terms differ only by the
// summation order of the reduction.
#include "traj.cuh"

namespace gm {

namespace {

constexpr int kTrajThreads = 128;

__device__ __forceinline__ double dsign(double v) { return (double)((v > 0.0) - (v < 0.0)); }

// scipy PchipInterpolator._edge_case: one-sided three-point estimate, shape preserving
__device__ __forceinline__ double pchip_edge(double h0, double h1, double m0, double m1) {
  const double num = __dsub_rn(__dmul_rn(__dadd_rn(__dmul_rn(2.0, h0), h1), m0), __dmul_rn(h0, m1));
  double d = __ddiv_rn(num, __dadd_rn(h0, h1));
  if (dsign(d) != dsign(m0)) return 0.0;
  if (dsign(m0) != dsign(m1) && fabs(d) > __dmul_rn(3.0, fabs(m0))) return __dmul_rn(3.0, m0);
  return d;
}

// y[0..nk) knot values, d[0..nk) <- pchip derivatives (uniform knot spacing h); threads of the block cooperate
__device__ void pchip_derivatives(const double* y, double* d, int nk, double h) {
  for (int k = threadIdx.x; k < nk; k += blockDim.x) {
    double out;
    if (nk == 2) {
      out = __ddiv_rn(__dsub_rn(y[1], y[0]), h);
    } else if (k == 0) {
      out = pchip_edge(h, h, __ddiv_rn(__dsub_rn(y[1], y[0]), h), __ddiv_rn(__dsub_rn(y[2], y[1]), h));
    } else if (k == nk - 1) {
      out = pchip_edge(h, h, __ddiv_rn(__dsub_rn(y[nk - 1], y[nk - 2]), h), __ddiv_rn(__dsub_rn(y[nk - 2], y[nk - 3]), h));
    } else {
      const double m0 = __ddiv_rn(__dsub_rn(y[k], y[k - 1]), h), m1 = __ddiv_rn(__dsub_rn(y[k + 1], y[k]), h);
      if (dsign(m0) != dsign(m1) || m0 == 0.0 || m1 == 0.0) {
        out = 0.0;
      } else {
        const double w1 = __dadd_rn(__dmul_rn(2.0, h), h), w2 = __dadd_rn(h, __dmul_rn(2.0, h));
        const double whmean = __ddiv_rn(__dadd_rn(__ddiv_rn(w1, m0), __ddiv_rn(w2, m1)), __dadd_rn(w1, w2));
        out = __ddiv_rn(1.0, whmean);
      }
    }
    d[k] = out;
  }
}

// CubicHermiteSpline coefficients + PPoly evaluation in the power basis (scipy _cubic.py / _ppoly evaluate_poly1)
__device__ __forceinline__ double pchip_eval(const double* y, const double* d, int nk, double h, int stride, int q) {
  int i = q / stride;
  if (i > nk - 2) i = nk - 2;
  const double s = (double)(q - i * stride);
  const double slope = __ddiv_rn(__dsub_rn(y[i + 1], y[i]), h);
  const double t = __ddiv_rn(__dsub_rn(__dadd_rn(d[i], d[i + 1]), __dmul_rn(2.0, slope)), h);
  const double c0 = __ddiv_rn(t, h);
  const double c1 = __dsub_rn(__ddiv_rn(__dsub_rn(slope, d[i]), h), t);
  double res = __dadd_rn(0.0, y[i]);
  double z = s;
  res = __dadd_rn(res, __dmul_rn(d[i], z));
  z = __dmul_rn(z, s);
  res = __dadd_rn(res, __dmul_rn(c1, z));
  z = __dmul_rn(z, s);
  res = __dadd_rn(res, __dmul_rn(c0, z));
  return res;
}

__device__ __forceinline__ double block_reduce(double v, double* red, int op /*0 sum, 1 max, 2 min*/) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const double u = __shfl_xor_sync(0xffffffffu, v, o);
    v = op == 0 ? v + u : (op == 1 ? fmax(v, u) : fmin(v, u));
  }
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  double r = red[0];
  for (int w = 1; w < (int)(blockDim.x >> 5); ++w) r = op == 0 ? r + red[w] : (op == 1 ? fmax(r, red[w]) : fmin(r, red[w]));
  return r;
}

__global__ void __launch_bounds__(kTrajThreads)
traj_interpolate_kernel(TrajParams p, const double* __restrict__ x, double* __restrict__ rot, double* __restrict__ ty,
                        double* __restrict__ terms) {
  __shared__ double s_y[2][kTrajMaxKnots];
  __shared__ double s_d[2][kTrajMaxKnots];
  __shared__ double s_red[kTrajThreads / 32];
  const int c = blockIdx.x, T = p.horizon, m = p.n_points;
  const double* xc = x + (size_t)c * 2 * m;
  double* rc = rot + (size_t)c * T;
  double* tc = ty + (size_t)c * T;
  double pen = 0.0;
  if (p.kind == 0) {
    // traj_utils.py:206-228: inc = clip(deg2rad(scale_rot * rad2deg(x)), +-max_rot); rot_{i+1} = rot_i + inc (sequential)
    const double r2d = 180.0 / 3.141592653589793238462643383279502884, d2r = 3.141592653589793238462643383279502884 / 180.0;
    if (threadIdx.x == 0) {
      double a = p.rx_init;
      rc[0] = a;
      for (int i = 0; i < m && i + 1 < T; ++i) {
        double inc = __dmul_rn(__dmul_rn(p.scale_rot, __dmul_rn(xc[i], r2d)), d2r);
        inc = fmin(fmax(inc, -p.max_rot), p.max_rot);
        a = __dadd_rn(a, inc);
        rc[i + 1] = a;
      }
    } else if (threadIdx.x == 32) {
      double a = 0.0;
      tc[0] = a;
      for (int i = 0; i < m && i + 1 < T; ++i) {
        double inc = __dmul_rn(p.scale_ty, xc[m + i]);
        inc = fmin(fmax(inc, -p.max_ty), p.max_ty);
        a = __dadd_rn(a, inc);
        tc[i + 1] = a;
      }
    }
  } else {
    // traj_utils.py:393-416: knots [rx_init, rx_init + x * scale_rot], [0, x * scale_ty] at 0, nr, 2 nr, ...
    const int nk = m + 1;
    const double h = (double)p.knot_stride;
    for (int k = threadIdx.x; k < nk; k += blockDim.x) {
      s_y[0][k] = k == 0 ? p.rx_init : __dadd_rn(p.rx_init, __dmul_rn(xc[k - 1], p.scale_rot));
      s_y[1][k] = k == 0 ? 0.0 : __dmul_rn(xc[m + k - 1], p.scale_ty);
    }
    __syncthreads();
    pchip_derivatives(s_y[0], s_d[0], nk, h);
    pchip_derivatives(s_y[1], s_d[1], nk, h);
    __syncthreads();
    for (int q = threadIdx.x; q < T; q += blockDim.x) {
      rc[q] = pchip_eval(s_y[0], s_d[0], nk, h, p.knot_stride, q);
      tc[q] = pchip_eval(s_y[1], s_d[1], nk, h, p.knot_stride, q);
    }
    if (terms != nullptr) {
      // compute_vel_noninterp (traj_utils.py:418-436): exp(max(|d knot| - max * nr)) over the parameter knots
      double worst = -1e300;
      for (int k = threadIdx.x; k + 1 < m; k += blockDim.x) {
        const double dr = __dsub_rn(__dmul_rn(xc[k + 1], p.scale_rot), __dmul_rn(xc[k], p.scale_rot));
        const double dt = __dsub_rn(__dmul_rn(xc[m + k + 1], p.scale_ty), __dmul_rn(xc[m + k], p.scale_ty));
        worst = fmax(worst, __dsub_rn(fabs(dr), __dmul_rn(p.max_rot, h)));
        worst = fmax(worst, __dsub_rn(fabs(dt), __dmul_rn(p.max_ty, h)));
      }
      worst = block_reduce(worst, s_red, 1);
      pen = exp(worst);
    }
  }
  if (terms == nullptr) return;
  __syncthreads();                         // this block's curve (global memory) is visible to all its threads
  // compute_vel_acc (traj_utils.py:161-165) + the normalised Frobenius norms
  double sv = 0.0, sa = 0.0, rmax = -1e300, rmin = 1e300;
  for (int i = threadIdx.x; i < T; i += blockDim.x) {
    const double r0 = rc[i], t0 = tc[i];
    rmax = fmax(rmax, r0);
    rmin = fmin(rmin, r0);
    if (i + 1 < T) {
      const double vr = __ddiv_rn(__dsub_rn(rc[i + 1], r0), p.vel_div[0]), vt = __ddiv_rn(__dsub_rn(tc[i + 1], t0), p.vel_div[1]);
      sv += __dadd_rn(__dmul_rn(vr, vr), __dmul_rn(vt, vt));
    }
    if (i + 2 < T) {
      const double ar = __ddiv_rn(__dadd_rn(__dsub_rn(rc[i + 2], __dmul_rn(2.0, rc[i + 1])), r0), p.acc_div[0]);
      const double at = __ddiv_rn(__dadd_rn(__dsub_rn(tc[i + 2], __dmul_rn(2.0, tc[i + 1])), t0), p.acc_div[1]);
      sa += __dadd_rn(__dmul_rn(ar, ar), __dmul_rn(at, at));
    }
  }
  sv = block_reduce(sv, s_red, 0);
  sa = block_reduce(sa, s_red, 0);
  if (p.kind == 0) {
    rmax = block_reduce(rmax, s_red, 1);
    rmin = block_reduce(rmin, s_red, 2);
    pen = (rmax > __dadd_rn(p.rx_init, p.rotation_limit) || rmin < __dsub_rn(p.rx_init, p.rotation_limit)) ? p.bound_penalty : 0.0;
  }
  if (threadIdx.x == 0) {
    terms[(size_t)c * 3 + 0] = sqrt(sv);
    terms[(size_t)c * 3 + 1] = sqrt(sa);
    terms[(size_t)c * 3 + 2] = pen;
  }
}

}  // namespace

int traj_interpolate(const TrajParams& p, const double* x, int n_cand, double* rot, double* ty, double* terms,
                     cudaStream_t stream) {
  GM_REQUIRE(p.kind == 0 || p.kind == 1, "traj_interpolate: kind=%d", p.kind);
  GM_REQUIRE(n_cand > 0 && p.horizon >= 2 && p.n_points >= 1, "traj_interpolate: bad sizes");
  if (p.kind == 0) {
    GM_REQUIRE(p.n_points == p.horizon - 1, "increments parametrisation: %d parameters per curve for horizon %d "
               "(the reference needs horizon == len(sample_traj) - 1, traj_utils.py:97-102)", p.n_points, p.horizon);
  } else {
    GM_REQUIRE(p.knot_stride >= 1 && p.n_points * p.knot_stride == p.horizon,
               "pchip parametrisation: %d knots x stride %d != horizon %d", p.n_points, p.knot_stride, p.horizon);
    GM_REQUIRE(p.n_points + 1 <= kTrajMaxKnots, "pchip parametrisation: more than %d knots", kTrajMaxKnots);
  }
  traj_interpolate_kernel<<<n_cand, kTrajThreads, 0, stream>>>(p, x, rot, ty, terms);
  GM_LAUNCHED(1);
  GM_LAUNCH_CHECK("traj_interpolate_kernel");
  return GM_OK;
}

}  // namespace gm
