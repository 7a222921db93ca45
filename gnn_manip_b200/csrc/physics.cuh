#pragma once
#include "common.cuh"

namespace gm {

// Dataset constants passed by value into kernels (schema: gnn_manip/utils/coffee_dataset.py:18-43).
struct PhysParams {
  int cart[3];   // cartesian_idx
  int mat;       // material_id
  int ctrl[3];   // control_idx (unused when has_ctrl == 0)
  int has_ctrl;
  int node_dim;  // 3*(k-1) + 6 + 1 (+3 with control)
  float vel_mean[3], vel_std[3], acc_mean[3], acc_std[3];
  float lower[3], upper[3];
  float conn_r;  // float32(conn_r): torch divides an f32 tensor by the f32-cast python scalar
};

struct RolloutParams {
  PhysParams phys;
  int k, D, n_per_graph, n_graphs;
  int n_rigid;              // rigid particles per graph
  const int* rigid_idx;     // [n_rigid] local node ids of the rigid particles
  const int* rigid_slot;    // [n_per_graph] inverse map (-1 = not rigid)
  int traj_steps;           // steps covered by the control trajectory
  const float* traj;        // optional precomputed [n_graphs, traj_steps, n_rigid, 3]; else rot/ty below
  const double* rot;        // [n_graphs, traj_steps] radians
  const double* ty;         // [n_graphs, traj_steps]
  const float* rigid_init;  // [n_rigid, 3] rigid positions of the initial state
  float ty_init[3];
  float* rigid_tgt;         // scratch [n_graphs, n_rigid, 3]
  int* step_dev;            // device step counter (advanced by integrate)
};

int node_features(const float* obs, int k, int n, int D, const PhysParams& p, float* nodes, cudaStream_t stream);
int edge_features(const float* pos, int stride, const int* send, const int* recv, const int* n_edges_dev,
                  int max_edges, float conn_r, float* edge_attr, cudaStream_t stream);
int apply_control(float* obs, const RolloutParams& rp, cudaStream_t stream);
int rigid_trajectory(const double* rot, const double* ty, int n_cand, int steps, const float* rigid_init, int n_rigid,
                     const float ty_init[3], float* out, cudaStream_t stream);
int random_walk_noise(const float* obs, int k, int n, int D, const PhysParams& p, const float* tgt, double noise_std,
                      unsigned long long seed, float* noisy_obs, float* noisy_tgt, float* samples, cudaStream_t stream);
int training_target(const float* obs, int k, int n, int D, const PhysParams& p, const float* tgt, float* target,
                    cudaStream_t stream);
int integrate(float* obs, const float* pred, const RolloutParams& rp, float* next_pos_out, cudaStream_t stream);

#ifdef __CUDACC__
// (p_send - p_recv)/r and its 2-norm (utils.py:52-59); the norm accumulates like torch.norm on CPU
// (x*x, then fused multiply-adds).
__device__ __forceinline__ float4 edge_feature(const float* __restrict__ pos, int stride, int s, int r,
                                               float conn_r) {
  const float* ps = pos + (size_t)s * stride;
  const float* pr = pos + (size_t)r * stride;
  const float x = __fdiv_rn(__fsub_rn(ps[0], pr[0]), conn_r);
  const float y = __fdiv_rn(__fsub_rn(ps[1], pr[1]), conn_r);
  const float z = __fdiv_rn(__fsub_rn(ps[2], pr[2]), conn_r);
  const float nn = __fsqrt_rn(__fmaf_rn(z, z, __fmaf_rn(y, y, __fmul_rn(x, x))));
  return make_float4(x, y, z, nn);
}
#endif

}  // namespace gm
