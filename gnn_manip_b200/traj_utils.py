"""Drop-in for gnn_manip/utils/traj_utils.py: CMAESolver / TrajectoryCMAsolver / InterpolatedCMAsolver with the
reference constructors and method names.  What changes underneath:

  * interpolate_trajectory -> one kernel over [candidates, steps] (gm_traj_interpolate) writing the control curves
    straight into the rollout plan's rot / ty buffers; compute_particles_tmatrix runs inside the plan;
  * cma_objective -> cma_objective_batch: the whole population (or this rank's contiguous slice of it) is rolled out
    as disjoint graphs of ONE device-resident plan, the Sinkhorn term of every candidate comes from one batched
    kernel sequence (gm_sinkhorn_divergence) reading the end states in place, and only [popsize] scalars leave the
    device (all_gather over the ranks, sharding.gather_population_values);
  * cma.fmin2 / cma.fmin_con -> cma_es.fmin_batch (host ask/tell, one batched objective call per generation).

There is no CPU implementation of any of the above; `device` must be a CUDA device.
"""
import ctypes

import numpy as np
import torch

from . import _lib, cma_es
from ._lib import check, gm_traj_param, ptr, stream_ptr
from .rollout_utils import Rollout
from .sharding import gather_population_values, partition_population


def sinkhorn_divergence(x, y, blur=0.05, scaling=0.5, diameter_max=None, x_idx=None):
    """geomloss.SamplesLoss("sinkhorn", p=2, blur)(x_b, y) for every cloud of a batch (uniform weights).
    x [B, N, 3] or [N, 3] (any float32 CUDA view with a contiguous last dimension), y [M, 3] -> [B] float32."""
    dev = _lib.compute_device(x)
    xx = x.to(device=dev, dtype=torch.float32)
    if xx.dim() == 2:
        xx = xx.unsqueeze(0)
    if xx.stride(-1) != 1:
        xx = xx.contiguous()
    yy = y.to(device=dev, dtype=torch.float32).contiguous()
    B, n_pts = xx.shape[0], xx.shape[1]
    idx = None
    if x_idx is not None:
        idx = x_idx.to(device=dev, dtype=torch.int32).contiguous()
        n_pts = idx.numel()
    if diameter_max is None:
        lo = torch.minimum(xx.reshape(-1, 3).min(0).values, yy.min(0).values)
        hi = torch.maximum(xx.reshape(-1, 3).max(0).values, yy.max(0).values)
        diameter_max = float((hi - lo).norm()) * 1.001 + 1e-12
    L = _lib.lib()
    ws_bytes = L.gm_sinkhorn_workspace_bytes(B, n_pts, yy.shape[0])
    ws = torch.empty((ws_bytes,), dtype=torch.uint8, device=dev)
    loss = torch.empty((B,), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        check(L.gm_sinkhorn_divergence(ptr(xx), xx.stride(1), xx.stride(0) if B > 1 else 0, ptr(idx), n_pts, ptr(yy),
                                       yy.shape[0], B, float(blur), float(scaling), float(diameter_max), ptr(loss),
                                       ptr(ws), ws_bytes, stream_ptr(dev)), "gm_sinkhorn_divergence")
    return loss


class CMAESolver:
    """traj_utils.py:14-194.  Same constructor; `model` is a gnn_manip_b200.EncProcDecGNN, `graph_attr` a
    gnn_manip_b200 GraphBoundedMultimaterialControl, `state_init` = (obs [k, N, D], next_pos)."""

    kind = None            # GM_TRAJ_* of the subclass

    def __init__(self, model, graph_attr, state_init, rx_init, ty_init, scale_rot, scale_ty, alpha, beta, gamma,
                 penalty, rho, device, cma_iter=10, cma_var=0.5, cma_popsize=21, cma_rand=1234, max_rot=1.9337,
                 max_ty=6.67e-4, total_steps=300, traj_points=10, impl=None, verbose=False):
        if verbose:
            print("Creating Trajectory Optimisation...")
        self.initial_state = state_init
        self.rx_init = np.deg2rad(rx_init)
        self.ty_init = ty_init
        self.sample_traj = None
        self.desired_pos = None
        self.model = model
        self.graph_attr = graph_attr
        self.horizon = total_steps
        self.device = torch.device(device) if not isinstance(device, torch.device) else device
        if self.device.type != "cuda":
            self.device = _lib.compute_device(None)           # raises without a GPU: there is no CPU path
        self.scale_rot = scale_rot
        self.nr_traj_points = traj_points
        self.traj_points = int(self.horizon / self.nr_traj_points)
        self.alpha, self.beta, self.gamma, self.penalty, self.rho = alpha, beta, gamma, penalty, rho
        self.max_rot = np.deg2rad(max_rot)
        self.max_ty = max_ty
        self.total_steps = total_steps
        self.left_limit, self.right_limit = 0.3, 0.7
        self.scale_ty = (self.ty_init[0] - self.left_limit) / self.scale_rot     # traj_utils.py:61 (overrides the arg)
        self.rotation_limit = 2.8973
        obs = state_init[0]
        mat = graph_attr.material_idx[0]
        self.rigid_particles_idx = obs[-1, :, mat] == 1
        self.coffee_particles_idx = obs[-1, :, mat] == 0
        self.rigid_particles = obs[-1, self.rigid_particles_idx, :][:, graph_attr.cartesian_idx]
        self.blur, self.sinkhorn_scaling = 0.05, 0.5          # geomloss.SamplesLoss("sinkhorn", p=2, blur=.05)
        self.cma_options = {"seed": cma_rand, "maxiter": cma_iter, "popsize": cma_popsize}
        self.cma_initial_var = cma_var
        self.impl = impl
        self._plans = {}
        self._sand_idx = self.coffee_particles_idx.nonzero().reshape(-1).to(device=self.device, dtype=torch.int32)
        lo = torch.as_tensor(graph_attr.bounds["lower_bounds"], dtype=torch.float32)
        hi = torch.as_tensor(graph_attr.bounds["upper_bounds"], dtype=torch.float32)
        self._diameter_max = float((hi - lo).norm()) * 2.0    # particles may leave the box a little; schedule bound only

    # ---- parametrisation (row N2) ------------------------------------------------------------------
    def _n_points(self):
        return self.sample_traj.shape[0]

    def traj_param(self):
        p = gm_traj_param()
        p.kind, p.horizon, p.n_points, p.knot_stride = self.kind, self.horizon, self._n_points(), self.nr_traj_points
        p.rx_init, p.scale_rot, p.scale_ty = float(self.rx_init), float(self.scale_rot), float(self.scale_ty)
        p.max_rot, p.max_ty = float(self.max_rot), float(self.max_ty)
        p.rotation_limit, p.bound_penalty = float(self.rotation_limit), 20.0
        p.vel_div[:], p.acc_div[:] = self._vel_div(), self._acc_div()
        return p

    def interpolate_population(self, X, rot=None, ty=None):
        """X [B, 2 * n_points] -> (rot [B, T], ty [B, T], terms [B, 3]) fp64 on the device."""
        X = np.atleast_2d(np.asarray(X, dtype=np.float64))
        B, T = X.shape[0], self.horizon
        if X.shape[1] != 2 * self._n_points():
            raise ValueError(f"search vector has {X.shape[1]} entries, the parametrisation needs {2 * self._n_points()}")
        dev = self.device
        xd = torch.from_numpy(np.ascontiguousarray(X)).to(dev)
        rot = torch.empty((B, T), dtype=torch.float64, device=dev) if rot is None else rot
        ty = torch.empty((B, T), dtype=torch.float64, device=dev) if ty is None else ty
        terms = torch.empty((B, 3), dtype=torch.float64, device=dev)
        p = self.traj_param()
        with torch.cuda.device(dev):
            check(_lib.lib().gm_traj_interpolate(ctypes.byref(p), ptr(xd), B, ptr(rot), ptr(ty), ptr(terms),
                                                 stream_ptr(dev)), "gm_traj_interpolate")
        return rot, ty, terms

    def interpolate_trajectory(self, x):
        """(traj_rot, traj_ty), each [horizon] (numpy, radians / metres)."""
        rot, ty, _ = self.interpolate_population(np.asarray(x, dtype=np.float64)[None])
        return rot[0].cpu().numpy(), ty[0].cpu().numpy()

    def get_rigid_body_trajectory_from_diff(self, x, demo=False):
        """traj_utils.py:87-103 -> (rigid_body_tensor [horizon, n_rigid, 3] on the CPU, actions [horizon, 2])."""
        dev = self.device
        if demo:
            x = np.asarray(x, dtype=np.float64)
            rot = torch.from_numpy(np.ascontiguousarray(x[: self.horizon, 0])).to(dev)[None]
            ty = torch.from_numpy(np.ascontiguousarray(x[: self.horizon, 1])).to(dev)[None]
        else:
            rot, ty, _ = self.interpolate_population(np.asarray(x, dtype=np.float64)[None])
        init = self.rigid_particles.to(device=dev, dtype=torch.float32).contiguous()
        out = torch.empty((self.horizon, init.shape[0], 3), dtype=torch.float32, device=dev)
        ty_init = (ctypes.c_float * 3)(*[float(v) for v in self.ty_init])
        with torch.cuda.device(dev):
            check(_lib.lib().gm_rigid_trajectory(ptr(rot), ptr(ty), 1, self.horizon, ptr(init), init.shape[0], ty_init,
                                                 ptr(out), stream_ptr(dev)), "gm_rigid_trajectory")
        actions = np.zeros((self.horizon, 2))
        actions[:, 0], actions[:, 1] = rot[0].cpu().numpy(), ty[0].cpu().numpy()
        return out.cpu(), actions

    def compute_particles_tmatrix(self, rotation, translation):
        """traj_utils.py:167-194 for one (rotation, translation) -> [n_rigid, 3]."""
        dev = self.device
        rot = torch.tensor([[float(rotation)]], dtype=torch.float64, device=dev)
        ty = torch.tensor([[float(translation)]], dtype=torch.float64, device=dev)
        init = self.rigid_particles.to(device=dev, dtype=torch.float32).contiguous()
        out = torch.empty((init.shape[0], 3), dtype=torch.float32, device=dev)
        ty_init = (ctypes.c_float * 3)(*[float(v) for v in self.ty_init])
        with torch.cuda.device(dev):
            check(_lib.lib().gm_rigid_trajectory(ptr(rot), ptr(ty), 1, 1, ptr(init), init.shape[0], ty_init, ptr(out),
                                                 stream_ptr(dev)), "gm_rigid_trajectory")
        return out.cpu()

    def compute_vel_acc(self, actions):
        vel = actions[1:, :] - actions[:-1, :]
        acc = actions[2:, :] - 2 * actions[1:-1, :] + actions[:-2, :]
        return vel, acc

    # ---- objective (rollout + row N1) --------------------------------------------------------------
    def _plan(self, n_graphs):
        plan = self._plans.get(n_graphs)
        if plan is None:
            zeros = np.zeros((n_graphs, self.horizon))
            plan = Rollout(self.model, self.graph_attr, self.initial_state[0], rot=zeros, ty=zeros,
                           ty_init=tuple(float(v) for v in self.ty_init), n_graphs=n_graphs, use_cuda_graph=True,
                           impl=self.impl, device=self.device)
            self._plans[n_graphs] = plan
        return plan

    def rollout_population(self, X):
        """Roll out every row of X [B, n_params] as a disjoint graph of one plan.
        Returns (plan, terms [B, 3] fp64 device): plan.obs holds the end states."""
        X = np.atleast_2d(np.asarray(X, dtype=np.float64))
        plan = self._plan(X.shape[0])
        _, _, terms = self.interpolate_population(X, plan.rot_dev, plan.ty_dev)
        plan.reset()
        plan.run(self.horizon)
        return plan, terms

    def wasserstein_population(self, plan):
        """Sinkhorn term of every candidate of `plan` against self.desired_pos, reading the sand rows of the last
        frame in place -> [B] float32 (device)."""
        if self.desired_pos is None:
            raise RuntimeError("desired_pos is not set (optimize_trajectory does it)")
        B, n, D = plan.n_graphs, plan.n_per_graph, plan.D
        cart0 = self.graph_attr.cartesian_idx[0]
        last = plan.obs[-1].reshape(B, n, D)[:, :, cart0: cart0 + 3]
        return sinkhorn_divergence(last, self.desired_pos, self.blur, self.sinkhorn_scaling,
                                   diameter_max=self._diameter_max, x_idx=self._sand_idx)

    def cma_objective_batch(self, X):
        """Objective of every row of X [popsize, n_params] -> numpy [popsize].  Under torch.distributed each rank rolls
        out its contiguous slice of the population (sharding.partition_population) and the values are all-gathered."""
        import torch.distributed as dist
        X = np.atleast_2d(np.asarray(X, dtype=np.float64))
        pop = X.shape[0]
        world, rank = (dist.get_world_size(), dist.get_rank()) if dist.is_available() and dist.is_initialized() else (1, 0)
        lo, hi = partition_population(pop, world, rank)
        if hi > lo:
            plan, terms = self.rollout_population(X[lo:hi])
            wass = self.wasserstein_population(plan).double()
            local = self.combine_loss(wass, terms)
        else:
            local = torch.zeros((0,), dtype=torch.float64, device=self.device)
        if world > 1:
            return gather_population_values(local, pop).cpu().numpy()
        return local.cpu().numpy()

    def cma_objective(self, x):
        """traj_utils.py:114-159 for one search vector."""
        return float(self.cma_objective_batch(np.asarray(x, dtype=np.float64)[None])[0])

    def combine_loss(self, wass, terms):
        raise NotImplementedError()

    def compute_loss(self, end_position, actions, cup_states=None, coffee_states=None, x=None):
        raise NotImplementedError()

    def set_sample_traj(self, sample_traj):
        raise NotImplementedError()

    def optimize_trajectory(self, desired_position):
        raise NotImplementedError()

    def _initial_params(self):
        m = self.sample_traj.shape[0]
        x0 = np.zeros(2 * m)
        x0[:m], x0[m:] = self.sample_traj[:, 0], self.sample_traj[:, 1]
        return x0

    def _wasserstein_single(self, end_position):
        return float(sinkhorn_divergence(end_position.to(self.device), self.desired_pos, self.blur,
                                         self.sinkhorn_scaling)[0])

    def close(self):
        for plan in self._plans.values():
            plan.close()
        self._plans = {}


class TrajectoryCMAsolver(CMAESolver):
    """traj_utils.py:197-285: per-step increments, horizon == len(sample_traj) - 1."""
    kind = _lib.GM_TRAJ_INCREMENTS

    def set_sample_traj(self, sample_traj):
        d = sample_traj[2:] - sample_traj[1:-1]
        self.sample_traj = np.stack((np.deg2rad(d[:, 0] / self.scale_rot), d[:, 1] / self.scale_ty)).T

    def _vel_div(self):
        return [float(self.max_rot), float(self.max_ty)]

    _acc_div = _vel_div

    def compute_vel_loss(self, vel):
        return float(np.linalg.norm(vel / np.array(self._vel_div())))

    def compute_acc_loss(self, acc):
        return float(np.linalg.norm(acc / np.array(self._acc_div())))

    def compute_boundaries_penalty(self, actions):
        r = actions[:, 0]
        if r.max() > self.rx_init + self.rotation_limit or r.min() < self.rx_init - self.rotation_limit:
            return 20.0
        return 0.0

    def combine_loss(self, wass, terms):
        return self.beta * wass + self.penalty * terms[:, 2] + self.alpha * terms[:, 0] + self.gamma * terms[:, 1]

    def compute_loss(self, end_position, actions, cup_states=None, coffee_states=None, x=None):
        w = self._wasserstein_single(end_position)
        vel, acc = self.compute_vel_acc(actions)
        v, a, b = self.compute_vel_loss(vel), self.compute_acc_loss(acc), self.compute_boundaries_penalty(actions)
        return self.beta * w + self.penalty * b + self.alpha * v + self.gamma * a, w, v, a, b, 0.0

    def optimize_trajectory(self, desired_position):
        self.desired_pos = desired_position.detach().to(device=self.device, dtype=torch.float32).contiguous()
        return cma_es.fmin_batch(self.cma_objective_batch, self._initial_params(), self.cma_initial_var,
                                 options=self.cma_options)


class InterpolatedCMAsolver(CMAESolver):
    """traj_utils.py:288-452: one knot every nr_traj_points steps, pchip in between."""
    kind = _lib.GM_TRAJ_PCHIP

    def set_sample_traj(self, sample_traj):
        idx = list(range(self.nr_traj_points, sample_traj.shape[0], self.nr_traj_points))
        pts = sample_traj[idx, :]
        self.sample_traj = np.stack(((np.deg2rad(pts[:, 0]) - self.rx_init) / self.scale_rot,
                                     (pts[:, 1] - self.ty_init[0]) / self.scale_ty)).T

    def _vel_div(self):
        return [1e-2, 4e-4]

    def _acc_div(self):
        return [2.2e-4, 1.45e-4]

    def compute_vel_loss(self, vel):
        return float(np.linalg.norm(vel / np.array(self._vel_div())))

    def compute_acc_loss(self, acc):
        return float(np.linalg.norm(acc / np.array(self._acc_div())))

    def compute_vel_noninterp(self, x):
        x = np.asarray(x, dtype=np.float64)
        rot, ty = x[: self.traj_points] * self.scale_rot, x[self.traj_points:] * self.scale_ty
        ineq = max((np.abs(rot[1:] - rot[:-1]) - self.max_rot * self.nr_traj_points).max(),
                   (np.abs(ty[1:] - ty[:-1]) - self.max_ty * self.nr_traj_points).max())
        return float(np.exp(ineq))

    def ineq_constraint(self, x):
        """traj_utils.py:375-391 for one vector [2 m] or a population [B, 2 m] -> [.., 2 m]."""
        X = np.atleast_2d(np.asarray(x, dtype=np.float64))
        tp = self.traj_points
        actions = np.zeros((X.shape[0], tp + 1, 2))
        actions[:, 1:, 0] = X[:, :tp] * self.scale_rot
        actions[:, 1:, 1] = X[:, tp:] * self.scale_ty
        vel = actions[:, 1:] - actions[:, :-1]
        up = np.abs(vel) - np.array([self.max_rot * self.nr_traj_points, self.max_ty * self.nr_traj_points])
        out = np.concatenate((up[:, :, 0] / self.scale_rot, up[:, :, 1] / self.scale_ty), axis=1)
        return out[0] if np.ndim(x) == 1 else out

    def combine_loss(self, wass, terms):
        return self.beta * wass + self.alpha * terms[:, 0] + self.gamma * terms[:, 1] + self.rho * terms[:, 2]

    def compute_loss(self, end_position, actions, cup_states=None, coffee_states=None, x=None):
        w = self._wasserstein_single(end_position)
        vel, acc = self.compute_vel_acc(actions)
        v, a = self.compute_vel_loss(vel), self.compute_acc_loss(acc)
        i = self.compute_vel_noninterp(x) if x is not None else 0.0
        return self.beta * w + self.alpha * v + self.gamma * a + self.rho * i, w, v, a, i, 0.0

    def optimize_trajectory(self, desired_position):
        self.desired_pos = desired_position.detach().to(device=self.device, dtype=torch.float32).contiguous()
        opts = dict(self.cma_options)
        opts["bounds"] = [-self.rotation_limit / self.scale_rot, self.rotation_limit / self.scale_rot]
        return cma_es.fmin_batch(self.cma_objective_batch, self._initial_params(), self.cma_initial_var, options=opts,
                                 constraints_batch=self.ineq_constraint)
