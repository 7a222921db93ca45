"""Drop-in for the hot-path parts of gnn_manip/utils/rollout_utils.py and the rollout loop of
gnn_manip/utils/traj_utils.py (CMAESolver.cma_objective, lines 119-152): the Euler integrator and a device-resident
rollout plan (graph build -> features -> 10 MP steps -> integrate, optionally replayed as one CUDA graph per step).
"""
import ctypes

import numpy as np
import torch

from . import _lib
from ._lib import check, gm_rollout_desc, ptr, stream_ptr


def get_position_from_prediction(stats, cartesian_idx, pred_acc, obs_seq):
    """gnn_manip/utils/rollout_utils.py:145-158: next_pos = p + ((p - p_prev) + (pred*acc_std + acc_mean))."""
    out_dev = obs_seq.device
    dev = _lib.compute_device(obs_seq)
    obs = obs_seq.to(device=dev, dtype=torch.float32).contiguous()
    pred = pred_acc.detach().to(device=dev, dtype=torch.float32).contiguous()
    k, n, D = obs.shape
    phys = _lib.make_phys(1.0, stats, None, cartesian_idx, 0, None)
    nxt = torch.empty((n, 3), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        check(_lib.lib().gm_next_position(ptr(obs), k, n, D, ctypes.byref(phys), ptr(pred), ptr(nxt),
                                          stream_ptr(dev)), "gm_next_position")
    return nxt.to(out_dev)


class Rollout:
    """Device-resident rollout of `n_graphs` independent candidates that share the initial state and the model
    (the loop body of traj_utils.py:123-152 == rollout_utils.py:38-65).

    graph_attr: a GraphBoundedMultimaterialControl (conn_r, stats, bounds, index lists, max_neighbours)
    obs:        initial state [k, N, D] (cloned; replicated per candidate)
    trajectory: rigid positions per step -- [T, n_rigid, 3] (shared) or [n_graphs, T, n_rigid, 3]; or
    rot, ty:    control curves [T] / [n_graphs, T] (radians, metres) fed to compute_particles_tmatrix on the device
    """

    def __init__(self, model, graph_attr, obs, trajectory=None, rot=None, ty=None, ty_init=(0.5, 0.5, 0.4),
                 n_graphs=1, use_cuda_graph=True, impl=None, device=None):
        L = _lib.lib()
        dev = torch.device(device) if device is not None else _lib.compute_device(obs)
        self.device = dev
        self.model = model
        self.packed = model.packed(dev)
        k, n, D = obs.shape
        self.k, self.n_per_graph, self.D, self.n_graphs = k, n, D, int(n_graphs)
        cart = list(graph_attr.cartesian_idx)
        if cart != [cart[0], cart[0] + 1, cart[0] + 2]:
            raise ValueError("the rollout kernels need contiguous cartesian_idx columns")
        mat = graph_attr.material_idx[0] if isinstance(graph_attr.material_idx, (list, tuple)) \
            else graph_attr.material_idx
        f32 = dict(device=dev, dtype=torch.float32)
        obs0 = obs.detach().to(**f32)
        self.obs0 = obs0.unsqueeze(1).expand(k, self.n_graphs, n, D).reshape(k, self.n_graphs * n, D).contiguous()
        self.obs = self.obs0.clone()
        rigid = (obs0[-1, :, mat] == 1).nonzero().reshape(-1).to(torch.int32)
        self.rigid_idx = rigid.contiguous()
        self.n_rigid = int(rigid.numel())
        self.pred = torch.zeros((self.n_graphs * n, 3), **f32)
        self._keep = []

        d = gm_rollout_desc()
        d.model = self.packed.handle
        d.impl = int(model.impl if impl is None else impl)
        d.phys = graph_attr.phys()
        d.k, d.D, d.n_per_graph, d.n_graphs = k, D, n, self.n_graphs
        # collate_utils.py:34: the noise-free path (the only one a rollout takes) calls get_connectivity without
        # max_neighbours, so the cap is utils.py:64's default 20 whatever the graph_attr was built with
        d.max_neighbours = int(graph_attr._simple_max_neighbours if graph_attr.noise_std is None
                               else graph_attr.max_neighbours)
        d.n_rigid = self.n_rigid
        d.rigid_idx = self.rigid_idx.data_ptr() if self.n_rigid else None
        d.traj_steps = 0
        if self.n_rigid and trajectory is not None:
            t = torch.as_tensor(trajectory).detach().to(**f32)
            if t.dim() == 3:
                t = t.unsqueeze(0).expand(self.n_graphs, *t.shape)
            if t.shape[0] != self.n_graphs or t.shape[2] != self.n_rigid or t.shape[3] != 3:
                raise ValueError(f"trajectory shape {tuple(t.shape)} does not match n_graphs/n_rigid")
            t = t.contiguous()
            self._keep.append(t)
            d.traj, d.traj_steps = t.data_ptr(), t.shape[1]
        elif self.n_rigid and rot is not None:
            r64 = torch.as_tensor(np.asarray(rot, dtype=np.float64)).reshape(-1, np.asarray(rot).shape[-1])
            t64 = torch.as_tensor(np.asarray(ty, dtype=np.float64)).reshape(-1, np.asarray(ty).shape[-1])
            if r64.shape[0] == 1:
                r64, t64 = r64.expand(self.n_graphs, -1), t64.expand(self.n_graphs, -1)
            if r64.shape != t64.shape or r64.shape[0] != self.n_graphs:
                raise ValueError("rot / ty must be [T] or [n_graphs, T]")
            r64, t64 = r64.contiguous().to(dev), t64.contiguous().to(dev)
            init = obs0[-1, rigid.long()][:, cart].contiguous()
            self._keep += [r64, t64, init]
            self.rot_dev, self.ty_dev, self.rigid_init = r64, t64, init      # [n_graphs, T] fp64 control curves
            d.rot, d.ty, d.rigid_init, d.traj_steps = r64.data_ptr(), t64.data_ptr(), init.data_ptr(), r64.shape[1]
        d.ty_init[:] = [float(x) for x in ty_init]
        d.obs = self.obs.data_ptr()
        d.pred = self.pred.data_ptr()
        d.use_cuda_graph = 1 if use_cuda_graph else 0
        self.desc = d
        self.traj_steps = int(d.traj_steps)

        ws_bytes = L.gm_rollout_workspace_bytes(ctypes.byref(d))
        if ws_bytes == 0:
            raise _lib.GmError("gm_rollout_workspace_bytes rejected the rollout description")
        self.workspace = torch.empty((ws_bytes,), dtype=torch.uint8, device=dev)
        self.handle = ctypes.c_void_p()
        with torch.cuda.device(dev):
            check(L.gm_rollout_create(ctypes.byref(self.handle), ctypes.byref(d), ptr(self.workspace), ws_bytes,
                                      stream_ptr(dev)), "gm_rollout_create")

    @property
    def launches_per_step(self):
        return int(_lib.load().gm_rollout_kernel_launches_per_step(self.handle))

    def reset(self, obs=None, step=0):
        """Restore the initial state (or `obs` [k, n_graphs*N, D]) and set the step counter."""
        with torch.cuda.device(self.device):
            self.obs.copy_(self.obs0 if obs is None else obs.reshape(self.obs.shape), non_blocking=True)
            check(_lib.lib().gm_rollout_reset(self.handle, int(step), stream_ptr(self.device)), "gm_rollout_reset")

    def run(self, steps):
        with torch.cuda.device(self.device):
            check(_lib.lib().gm_rollout_run(self.handle, int(steps), stream_ptr(self.device)), "gm_rollout_run")
        return self.obs

    def graph_handles(self):
        """Live views into the plan's workspace: (nbr [n,K], cnt [n], edge_ptr [n+1], n_edges [1], nodes [n,F]) of the
        most recent step (overwritten by the next one)."""
        L = _lib.lib()
        p = [ctypes.c_void_p() for _ in range(5)]
        check(L.gm_rollout_graph(self.handle, *[ctypes.byref(x) for x in p]), "gm_rollout_graph")
        n, K, F = self.n_graphs * self.n_per_graph, self.desc.max_neighbours, self.model.node_dim
        base = self.workspace.data_ptr()

        def view(pv, count, dtype):
            off = pv.value - base
            return self.workspace[off: off + count * 4].view(dtype)
        return (view(p[0], n * K, torch.int32).reshape(n, K), view(p[1], n, torch.int32),
                view(p[2], n + 1, torch.int32), view(p[3], 1, torch.int32),
                view(p[4], n * F, torch.float32).reshape(n, F))

    def graph(self):
        """Copies of graph_handles()."""
        return tuple(t.clone() for t in self.graph_handles())

    def close(self):
        if getattr(self, "handle", None):
            _lib.load().gm_rollout_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def rollout_final_state(model, graph_attr, state, trajectory, steps, **kw):
    """Device-resident rollout of `steps` steps from `state` under a rigid-body trajectory [T, n_rigid, 3]: the final
    observation window [k, N, D] on the device of the input state (the end state cma_objective scores)."""
    obs = state[0] if isinstance(state, (tuple, list)) else state
    plan = Rollout(model, graph_attr, obs, trajectory=trajectory, **kw)
    out = plan.run(steps).to(obs.device).clone()
    plan.close()
    return out


def perform_prediction(dataset, model, current_state, device=None):
    """gnn_manip/utils/rollout_utils.py:95-107: one prediction through the reference-shaped calls (graph + features,
    model forward, Euler step) -- every one of them a CUDA kernel sequence behind the C ABI."""
    batch = dataset.graph_attr.process_collate([current_state])
    pred_acc = model.forward(batch[0], batch[1], batch[2])
    return get_position_from_prediction(dataset.stats, dataset.cartesian_idx, pred_acc, current_state[0])


def compute_rollout(dataset, model, args):
    """gnn_manip/utils/rollout_utils.py:14-67, same contract: returns prediction [T, N, D] (numpy float64), row i = the
    last frame of the state after the rigid control of step i was written and before the model moved it.

    args.cma_traj (path of a saved [T, 2] (rotation [rad], translation) array, :24-31): the whole loop runs as one
    device-resident plan (CUDA-graph replay per step); only the recorded frames come back.  Without it the rigid
    particles follow the dataset's ground truth (:33, 45, 60) through the per-step reference-shaped calls.
    Unlike the reference, dataset[0]'s stored tensor is NOT mutated (the reference advances it in place, :37, 47-54)."""
    k_steps = getattr(args, "k_steps", dataset.k)
    cma_traj = getattr(args, "cma_traj", None)
    nof_steps = dataset.time_steps if cma_traj is not None else dataset.time_steps - k_steps
    obs0 = dataset[0][0]
    nof_particles, data_dim = obs0.shape[1], obs0.shape[2]
    rigid = obs0[0, :, dataset.material_id] == 1
    prediction = np.zeros((nof_steps, nof_particles, data_dim))
    if cma_traj is not None:
        actions = np.load(cma_traj) if isinstance(cma_traj, str) else np.asarray(cma_traj)
        ty_init = getattr(args, "ty_init", (0.5, 0.5, 0.4))
        plan = Rollout(model, dataset.graph_attr, obs0, rot=actions[:nof_steps, 0], ty=actions[:nof_steps, 1],
                       ty_init=tuple(ty_init), use_cuda_graph=True,
                       device=getattr(args, "device", None) if str(getattr(args, "device", "cpu")) != "cpu" else None)
        frames = torch.empty((nof_steps, nof_particles, data_dim), dtype=torch.float32, device=plan.device)
        for i in range(nof_steps):
            plan.run(1)
            frames[i].copy_(plan.obs[-2], non_blocking=True)     # the pre-shift last frame, control columns included
        prediction[:] = frames.cpu().numpy()
        plan.close()
        return prediction
    groundtruth = torch.stack([dataset[i][0][-1] for i in range(nof_steps)]).float()
    current = (obs0.clone(), dataset[0][1].clone())
    cart, ctrl = dataset.cartesian_idx, dataset.control_idx
    for i in range(nof_steps):
        new_rigid = current[0][-1, rigid, :].clone()
        new_rigid[:, ctrl] = groundtruth[i, rigid, :][:, ctrl]
        current[0][-1, rigid, :] = new_rigid
        prediction[i] = current[0][-1].numpy()
        next_pos = perform_prediction(dataset, model, current)
        current[0][:-1] = current[0][1:].clone()
        current[0][-1][:, cart] = next_pos.detach().clone()
        new_rigid[:, cart] = groundtruth[i, rigid, :][:, cart]
        current[0][-1, rigid, :] = new_rigid.clone()
    return prediction
