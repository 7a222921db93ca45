// Rows A2 and A5 -- node/edge feature build and the Euler integrator / rollout bookkeeping.
//
// A2 replaces GraphBoundedMultimaterial(Control).compute_nodes (gnn_manip/utils/collate_utils.py:195-232),
// get_nodes_vel (gnn_manip/utils/utils.py:27-40) and get_edges_displacement (utils.py:43-61).
// A5 replaces get_position_from_prediction (gnn_manip/utils/rollout_utils.py:145-158), the state
// bookkeeping of CMAESolver.cma_objective (gnn_manip/utils/traj_utils.py:119-152) and
// compute_particles_tmatrix (traj_utils.py:167-194).
//
// All arithmetic is IEEE fp32 with the reference's operation order and *no* FMA contraction
// (true division, sqrt.rn), so node features and the integrator are bit-identical to the CPU
// reference; the edge norm uses the fused accumulation torch.norm performs on CPU.
#include <string.h>
#include <math.h>
#include <curand_kernel.h>
#include "common.cuh"
#include "physics.cuh"

namespace gm {

namespace {

__global__ void node_features_kernel(const float* __restrict__ obs, int k, int n, int D, PhysParams p,
                                     float* __restrict__ nodes) {
  pdl_wait();                                 // programmatic dependent launch: see common.cuh
  pdl_launch_dependents();
  const int F = p.node_dim;
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (long long)n * F) return;
  const int node = (int)(idx / F), f = (int)(idx - (long long)node * F);
  const int nvel = 3 * (k - 1);
  const float* last = obs + ((size_t)(k - 1) * n + node) * D;
  float out;
  if (f < nvel) {
    const int t = f / 3, d = f - 3 * t;
    const float a = obs[((size_t)(t + 1) * n + node) * D + p.cart[d]];
    const float b = obs[((size_t)t * n + node) * D + p.cart[d]];
    out = __fdiv_rn(__fsub_rn(__fsub_rn(a, b), p.vel_mean[d]), p.vel_std[d]);
  } else if (f < nvel + 6) {
    const int b = f - nvel, d = b % 3;
    const float x = last[p.cart[d]];
    const float dist = b < 3 ? __fsub_rn(x, p.lower[d]) : __fsub_rn(p.upper[d], x);
    const float q = __fdiv_rn(dist, p.conn_r);
    out = fminf(fmaxf(q, -1.0f), 1.0f);
    if (q != q) out = q;  // torch.clamp propagates NaN
  } else if (f == nvel + 6) {
    out = last[p.mat];
  } else {
    const int d = f - (nvel + 7);
    out = __fdiv_rn(__fsub_rn(last[p.ctrl[d]], p.vel_mean[d]), p.vel_std[d]);
  }
  nodes[idx] = out;
}

__global__ void edge_features_kernel(const float* __restrict__ pos, int stride, const int* __restrict__ send,
                                     const int* __restrict__ recv, const int* __restrict__ n_edges_dev,
                                     int max_edges, float conn_r, float4* __restrict__ edge_attr) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  const int ne = n_edges_dev ? min(*n_edges_dev, max_edges) : max_edges;
  if (e >= ne) return;
  edge_attr[e] = edge_feature(pos, stride, send[e], recv[e], conn_r);
}

// Rigid ("cup") particle r of graph g at rollout step i: rotation about X by rot, Y shift by ty around
// ty_init with y/z swapped and the offset taken as (ty_init - p) (traj_utils.py:167-194).
__device__ __forceinline__ void rigid_target(const RolloutParams& rp, int g, int r, int step, float out[3]) {
  if (rp.traj != nullptr) {
    const float* t = rp.traj + (((size_t)g * rp.traj_steps + step) * rp.n_rigid + r) * 3;
    out[0] = t[0]; out[1] = t[1]; out[2] = t[2];
    return;
  }
  const double rot = rp.rot[(size_t)g * rp.traj_steps + step];
  const double ty = rp.ty[(size_t)g * rp.traj_steps + step];
  const float c = (float)cos(rot), s = (float)sin(rot);
  const float t1 = (float)((double)rp.ty_init[1] + ty);   // torch.tensor(..., dtype=float32) of an f64 sum
  const float* p0 = rp.rigid_init + (size_t)r * 3;
  const float q0 = __fsub_rn(rp.ty_init[0], p0[0]);
  const float q1 = __fsub_rn(rp.ty_init[1], p0[2]);
  const float q2 = __fsub_rn(rp.ty_init[2], p0[1]);
  const float w0 = __fadd_rn(q0, rp.ty_init[0]);
  const float w1 = __fadd_rn(__fadd_rn(__fmul_rn(c, q1), __fmul_rn(-s, q2)), t1);
  const float w2 = __fadd_rn(__fadd_rn(__fmul_rn(s, q1), __fmul_rn(c, q2)), rp.ty_init[2]);
  out[0] = w0; out[1] = w2; out[2] = w1;
}

// get_rigid_body_trajectory_from_diff (traj_utils.py:87-103): the rigid-particle targets of every candidate and step,
// materialised (the rollout plan itself evaluates rigid_target on the fly; this is the reference-shaped output).
__global__ void rigid_trajectory_kernel(RolloutParams rp, int steps, float* __restrict__ out) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long total = (long long)rp.n_graphs * steps * rp.n_rigid;
  if (idx >= total) return;
  const int r = (int)(idx % rp.n_rigid);
  const int step = (int)((idx / rp.n_rigid) % steps);
  const int g = (int)(idx / ((long long)rp.n_rigid * steps));
  float tgt[3];
  rigid_target(rp, g, r, step, tgt);
  out[idx * 3 + 0] = tgt[0]; out[idx * 3 + 1] = tgt[1]; out[idx * 3 + 2] = tgt[2];
}

// Row N3: random_walk_noise (utils.py:96-115) + the noisy observation / target of _process_noisy (collate_utils.py:169-193).
// One thread per (particle, axis): k-1 normal draws, two running sums.
__global__ void random_walk_noise_kernel(const float* __restrict__ obs, int k, int n, int D, PhysParams p,
                                         const float* __restrict__ tgt, float step_std, unsigned long long seed,
                                         float* __restrict__ noisy_obs, float* __restrict__ noisy_tgt,
                                         float* __restrict__ samples) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n * 3) return;
  const int node = idx / 3, d = idx - 3 * node;
  curandStatePhilox4_32_10_t st;
  curand_init(seed, (unsigned long long)idx, 0ull, &st);
  float vel = 0.f, posn = 0.f;
  // frame 0 carries no noise (utils.py:111-113)
  for (int t = 0; t < k; ++t) {
    if (t > 0) {
      const float z = curand_normal(&st) * step_std;
      if (samples) samples[((size_t)(t - 1) * n + node) * 3 + d] = z;
      vel = __fadd_rn(vel, z);              // torch.cumsum over the velocity noise
      posn = __fadd_rn(posn, vel);          // torch.cumsum over the noisy velocity
    }
    const size_t o = ((size_t)t * n + node) * D + p.cart[d];
    noisy_obs[o] = __fadd_rn(obs[o], posn);
  }
  if (noisy_tgt) noisy_tgt[idx] = __fadd_rn(tgt[idx], posn);
}

__global__ void copy_non_cartesian_kernel(const float* __restrict__ obs, long long total, int D, PhysParams p,
                                          float* __restrict__ out) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const int c = (int)(i % D);
  if (c != p.cart[0] && c != p.cart[1] && c != p.cart[2]) out[i] = obs[i];
}

// compute_target (collate_utils.py:148-157) over compute_acceleration (utils.py:10-24)
__global__ void training_target_kernel(const float* __restrict__ obs, int k, int n, int D, PhysParams p,
                                       const float* __restrict__ tgt, float* __restrict__ target) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n * 3) return;
  const int node = idx / 3, d = idx - 3 * node;
  const float p1 = obs[((size_t)(k - 1) * n + node) * D + p.cart[d]];
  const float p2 = obs[((size_t)(k - 2) * n + node) * D + p.cart[d]];
  const float acc = __fadd_rn(__fsub_rn(tgt[idx], __fmul_rn(2.0f, p1)), p2);
  target[idx] = __fdiv_rn(__fsub_rn(acc, p.acc_mean[d]), p.acc_std[d]);
}

// Step (1) of the loop body: control columns of the rigid particles (traj_utils.py:126-135).
__global__ void apply_control_kernel(float* __restrict__ obs, RolloutParams rp) {
  pdl_wait();                                 // programmatic dependent launch: see common.cuh
  pdl_launch_dependents();
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= rp.n_graphs * rp.n_rigid) return;
  const int g = idx / rp.n_rigid, r = idx - g * rp.n_rigid;
  const int step = *rp.step_dev;
  const int node = g * rp.n_per_graph + rp.rigid_idx[r];
  const int n = rp.n_graphs * rp.n_per_graph;
  float* last = obs + ((size_t)(rp.k - 1) * n + node) * rp.D;
  float tgt[3];
  if (step < rp.traj_steps) {
    rigid_target(rp, g, r, step, tgt);
#pragma unroll
    for (int d = 0; d < 3; ++d) {
      rp.rigid_tgt[(size_t)idx * 3 + d] = tgt[d];
      last[rp.phys.ctrl[d]] = __fsub_rn(tgt[d], last[rp.phys.cart[d]]);
    }
  } else {  // reference quirk traj_utils.py:128-129: control := position, state left to the model
#pragma unroll
    for (int d = 0; d < 3; ++d) last[rp.phys.ctrl[d]] = last[rp.phys.cart[d]];
  }
}

// Steps (3)+(4): semi-implicit Euler, window shift, rigid overwrite (rollout_utils.py:145-158,
// traj_utils.py:146-152). One thread per node; each node's rows are private to its thread.
__global__ void integrate_kernel(float* __restrict__ obs, const float* __restrict__ pred, RolloutParams rp,
                                 float* __restrict__ next_pos_out) {
  pdl_wait();                                 // programmatic dependent launch: see common.cuh
  pdl_launch_dependents();
  const int node = blockIdx.x * blockDim.x + threadIdx.x;
  const int n = rp.n_graphs * rp.n_per_graph;
  const int k = rp.k, D = rp.D;
  if (node < n) {
  const PhysParams& p = rp.phys;
  float nxt[3];
  {
    const float* last = obs + ((size_t)(k - 1) * n + node) * D;
    const float* prev = obs + ((size_t)(k - 2) * n + node) * D;
#pragma unroll
    for (int d = 0; d < 3; ++d) {
      const float acc = __fadd_rn(__fmul_rn(pred[(size_t)node * 3 + d], p.acc_std[d]), p.acc_mean[d]);
      const float lp = last[p.cart[d]];
      const float vel = __fadd_rn(__fsub_rn(lp, prev[p.cart[d]]), acc);
      nxt[d] = __fadd_rn(lp, vel);
    }
  }
  if (next_pos_out) {
#pragma unroll
    for (int d = 0; d < 3; ++d) next_pos_out[(size_t)node * 3 + d] = nxt[d];
  }
  for (int t = 0; t + 1 < k; ++t) {
    const float* src = obs + ((size_t)(t + 1) * n + node) * D;
    float* dst = obs + ((size_t)t * n + node) * D;
    for (int c = 0; c < D; ++c) dst[c] = src[c];
  }
  float* last = obs + ((size_t)(k - 1) * n + node) * D;
  const int local = node % rp.n_per_graph;
  const int rslot = rp.rigid_slot ? rp.rigid_slot[local] : -1;
  const int step = *rp.step_dev;
  if (rslot >= 0) {
    // traj_utils.py:149-152: the whole pre-shift rigid row is written back; its position columns are the control
    // target while the trajectory lasts and the *old* position afterwards (rigid particles then stay frozen).
    if (step < rp.traj_steps) {
      const int g = node / rp.n_per_graph;
      const float* tgt = rp.rigid_tgt + ((size_t)g * rp.n_rigid + rslot) * 3;
#pragma unroll
      for (int d = 0; d < 3; ++d) last[p.cart[d]] = tgt[d];
    }
  } else {
#pragma unroll
    for (int d = 0; d < 3; ++d) last[p.cart[d]] = nxt[d];
  }
  }
  // step counter: the last block to get here advances it (every block has read it by then); step_dev[1] is the ticket
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const int t = atomicAdd(rp.step_dev + 1, 1);
    if (t == (int)gridDim.x - 1) {
      rp.step_dev[1] = 0;
      rp.step_dev[0] += 1;
    }
  }
}

}  // namespace

int node_features(const float* obs, int k, int n, int D, const PhysParams& p, float* nodes, cudaStream_t stream) {
  GM_REQUIRE(k >= 2 && n > 0 && D > 0, "node_features: bad shape k=%d n=%d D=%d", k, n, D);
  const long long total = (long long)n * p.node_dim;
  GM_CUDA(launch_pdl(node_features_kernel, (int)ceil_div(total, 256), 256, 0, stream, 1, obs, k, n, D, p, nodes));
  GM_LAUNCHED(1);
  GM_LAUNCH_CHECK("node_features_kernel");
  return GM_OK;
}

int edge_features(const float* pos, int stride, const int* send, const int* recv, const int* n_edges_dev,
                  int max_edges, float conn_r, float* edge_attr, cudaStream_t stream) {
  if (max_edges <= 0) return GM_OK;
  edge_features_kernel<<<ceil_div(max_edges, 256), 256, 0, stream>>>(pos, stride, send, recv, n_edges_dev,
                                                                    max_edges, conn_r, (float4*)edge_attr);
  GM_LAUNCHED(1);
  GM_LAUNCH_CHECK("edge_features_kernel");
  return GM_OK;
}

int rigid_trajectory(const double* rot, const double* ty, int n_cand, int steps, const float* rigid_init, int n_rigid,
                     const float ty_init[3], float* out, cudaStream_t stream) {
  GM_REQUIRE(rot != nullptr && ty != nullptr && rigid_init != nullptr && out != nullptr, "rigid_trajectory: NULL argument");
  GM_REQUIRE(n_cand > 0 && steps > 0 && n_rigid > 0, "rigid_trajectory: bad sizes");
  RolloutParams rp;
  memset(&rp, 0, sizeof(rp));
  rp.n_graphs = n_cand; rp.n_rigid = n_rigid; rp.traj_steps = steps;
  rp.rot = rot; rp.ty = ty; rp.rigid_init = rigid_init;
  for (int i = 0; i < 3; ++i) rp.ty_init[i] = ty_init[i];
  const long long total = (long long)n_cand * steps * n_rigid;
  rigid_trajectory_kernel<<<ceil_div(total, 256), 256, 0, stream>>>(rp, steps, out);
  GM_LAUNCHED(1);
  GM_LAUNCH_CHECK("rigid_trajectory_kernel");
  return GM_OK;
}

int random_walk_noise(const float* obs, int k, int n, int D, const PhysParams& p, const float* tgt, double noise_std,
                      unsigned long long seed, float* noisy_obs, float* noisy_tgt, float* samples, cudaStream_t stream) {
  GM_REQUIRE(obs != nullptr && noisy_obs != nullptr && k >= 2 && n > 0 && D >= 3, "random_walk_noise: bad arguments");
  GM_REQUIRE(noisy_tgt == nullptr || tgt != nullptr, "random_walk_noise: noisy_tgt needs tgt");
  const long long total = (long long)k * n * D;
  if (noisy_obs != obs) {
    copy_non_cartesian_kernel<<<ceil_div(total, 256), 256, 0, stream>>>(obs, total, D, p, noisy_obs);
    GM_LAUNCHED(1);
  }
  const float step_std = (float)(noise_std / sqrt((double)(k - 1)));     // utils.py:103
  random_walk_noise_kernel<<<ceil_div((long long)n * 3, 128), 128, 0, stream>>>(obs, k, n, D, p, tgt, step_std, seed,
                                                                            noisy_obs, noisy_tgt, samples);
  GM_LAUNCHED(1);
  GM_LAUNCH_CHECK("random_walk_noise kernels");
  return GM_OK;
}

int training_target(const float* obs, int k, int n, int D, const PhysParams& p, const float* tgt, float* target,
                    cudaStream_t stream) {
  GM_REQUIRE(obs != nullptr && tgt != nullptr && target != nullptr && k >= 2 && n > 0, "training_target: bad arguments");
  training_target_kernel<<<ceil_div((long long)n * 3, 256), 256, 0, stream>>>(obs, k, n, D, p, tgt, target);
  GM_LAUNCHED(1);
  GM_LAUNCH_CHECK("training_target_kernel");
  return GM_OK;
}

int apply_control(float* obs, const RolloutParams& rp, cudaStream_t stream) {
  const int total = rp.n_graphs * rp.n_rigid;
  if (total <= 0) return GM_OK;
  GM_CUDA(launch_pdl(apply_control_kernel, (int)ceil_div(total, 128), 128, 0, stream, 1, obs, rp));
  GM_LAUNCHED(1);
  GM_LAUNCH_CHECK("apply_control_kernel");
  return GM_OK;
}

int integrate(float* obs, const float* pred, const RolloutParams& rp, float* next_pos_out, cudaStream_t stream) {
  const int n = rp.n_graphs * rp.n_per_graph;
  GM_CUDA(launch_pdl(integrate_kernel, ceil_div(n, 128), 128, 0, stream, 1, obs, pred, rp, next_pos_out));
  GM_LAUNCHED(1);
  GM_LAUNCH_CHECK("integrate_kernel");
  return GM_OK;
}

GM_TRACE_DUMP(trace_dump_physics)

}  // namespace gm
