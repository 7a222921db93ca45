#pragma once
#include "common.cuh"

namespace gm {

constexpr int kTrajMaxKnots = 1024;   // pchip knots (first one included) held in shared memory per candidate

// Trajectory parametrisation of the CMA-ES solvers (gnn_manip/utils/traj_utils.py); all values fp64 like numpy.
struct TrajParams {
  int kind;            // 0: increments (TrajectoryCMAsolver), 1: pchip knots (InterpolatedCMAsolver)
  int horizon;         // control steps produced per candidate
  int n_points;        // parameters per curve: increments: horizon - 1; pchip: knots after the first
  int knot_stride;     // pchip: steps between knots (nr_traj_points)
  double rx_init;      // initial rotation, radians
  double scale_rot, scale_ty;
  double max_rot, max_ty;        // radians / metres per step
  double rotation_limit;         // increments: bound penalty limit (traj_utils.py:64)
  double bound_penalty;          // increments: 20.0 (traj_utils.py:268-272)
  double vel_div[2], acc_div[2]; // normalisers of the velocity / acceleration terms
};

// x [n_cand, 2 * n_points] -> rot, ty [n_cand, horizon]; terms (optional) [n_cand, 3] = velocity term, acceleration
// term, penalty (bound penalty for kind 0, exp(max inequality) of the knots for kind 1).
int traj_interpolate(const TrajParams& p, const double* x, int n_cand, double* rot, double* ty, double* terms,
                     cudaStream_t stream);

}  // namespace gm
