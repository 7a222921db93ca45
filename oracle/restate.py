"""CPU restatement of the gnn-manip rollout hot path.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py): the checker for tests/, smoke() and the
``cpu_baseline`` / ``--impl reference`` legs of bench.py.  Never imported by gnn_manip_b200.

Parity: PINNED against the unmodified reference (tests/golden/*.npz written by
oracle/make_golden.py, and live in tests/test_oracle_vs_reference.py whenever /root/reference
is present) and against scikit-learn's KDTree (the reference's third-party graph builder).

Every function cites the reference lines it restates (paths relative to /root/reference).
"""
import ctypes
import math
import os

import numpy as np
import torch
from torch import nn

_HERE = os.path.dirname(os.path.abspath(__file__))

# --------------------------------------------------------------------------------------
# A1  radius graph
# --------------------------------------------------------------------------------------


def get_connectivity_sklearn(pos_nodes, conn_r, max_neighbours=20):
    """The reference's own third-party call sequence (gnn_manip/utils/utils.py:64-93):
    KDTree.query_radius(sorted by distance) then keep the first ``max_neighbours`` per query.
    Used to pin the restatements below and as the graph phase of the CPU baseline."""
    from sklearn.neighbors import KDTree
    pts = np.asarray(pos_nodes)
    tree = KDTree(pts)
    nbrs, _ = tree.query_radius(pts, conn_r, return_distance=True, sort_results=True)
    kept = [r[:max_neighbours] for r in nbrs]
    counts = np.fromiter((len(r) for r in kept), dtype=np.int64, count=len(kept))
    senders = np.repeat(np.arange(len(kept), dtype=np.int64), counts)
    receivers = np.concatenate(kept).astype(np.int64) if len(kept) else np.zeros(0, np.int64)
    return torch.from_numpy(senders), torch.from_numpy(receivers)


def get_connectivity_bruteforce(pos_nodes, conn_r, max_neighbours=20):
    """O(N^2) numpy restatement of utils.py:64-93 + sklearn KDTree64 arithmetic:
    inputs cast f32->f64 (sklearn/neighbors/_binary_tree.pxi.tp:1267), d2 = ((dx*dx)+dy*dy)+dz*dz
    accumulated in f64 (metrics/_dist_metrics.pxd.tp:39-49), keep d2 <= r*r
    (_binary_tree.pxi.tp:1952-1967), order by ascending distance (ties: by index -- the
    reference leaves ties undefined, utils/_sorting.pyx:63-108), keep first max_neighbours
    (utils.py:80-85).  Query node = sender, neighbour = receiver, self-loop included."""
    p = np.asarray(pos_nodes, dtype=np.float32).astype(np.float64)
    n = p.shape[0]
    r2 = float(conn_r) * float(conn_r)
    send, recv = [], []
    for i in range(n):
        d = p - p[i]
        d2 = d[:, 0] * d[:, 0]
        d2 = d2 + d[:, 1] * d[:, 1]
        d2 = d2 + d[:, 2] * d[:, 2]
        idx = np.nonzero(d2 <= r2)[0]
        order = np.lexsort((idx, d2[idx]))[:max_neighbours]
        recv.append(idx[order])
        send.append(np.full(len(order), i, dtype=np.int64))
    s = np.concatenate(send) if send else np.zeros(0, np.int64)
    r = np.concatenate(recv).astype(np.int64) if recv else np.zeros(0, np.int64)
    return torch.from_numpy(s), torch.from_numpy(r)


_clib = None


def _load_clib():
    global _clib
    if _clib is None:
        path = os.path.join(_HERE, "libradius_graph_oracle.so")
        if not os.path.exists(path):
            raise RuntimeError(f"{path} missing: run `make -C oracle` (or __graft_entry__.build())")
        lib = ctypes.CDLL(path)
        lib.oracle_radius_graph.restype = ctypes.c_int
        lib.oracle_radius_graph.argtypes = [
            ctypes.c_void_p, ctypes.c_int, ctypes.c_double, ctypes.c_int,
            ctypes.c_void_p, ctypes.c_void_p]
        _clib = lib
    return _clib


def neighbour_table_c(pos_nodes, conn_r, max_neighbours=20):
    """Cell-grid C restatement (oracle/radius_graph.c), same rule as the brute force above.
    Returns (nbr [N,K] i32 padded with -1, cnt [N] i32)."""
    p = np.ascontiguousarray(np.asarray(pos_nodes, dtype=np.float32))
    n = p.shape[0]
    nbr = np.full((n, max_neighbours), -1, dtype=np.int32)
    cnt = np.zeros(n, dtype=np.int32)
    rc = _load_clib().oracle_radius_graph(p.ctypes.data, n, float(conn_r), int(max_neighbours),
                                           nbr.ctypes.data, cnt.ctypes.data)
    if rc != 0:
        raise RuntimeError(f"oracle_radius_graph failed rc={rc}")
    return nbr, cnt


def table_to_edges(nbr, cnt):
    """(nbr, cnt) -> the reference's sender-major / distance-minor edge list (utils.py:87-91)."""
    nbr = np.asarray(nbr)
    cnt = np.asarray(cnt).astype(np.int64)
    n, k = nbr.shape
    mask = np.arange(k)[None, :] < cnt[:, None]
    senders = np.repeat(np.arange(n, dtype=np.int64), cnt)
    receivers = nbr[mask].astype(np.int64)
    return torch.from_numpy(senders), torch.from_numpy(receivers)


def get_connectivity(pos_nodes, conn_r, max_neighbours=20):
    """Restated get_connectivity (utils.py:64-93) via the C cell grid."""
    nbr, cnt = neighbour_table_c(pos_nodes, conn_r, max_neighbours)
    return table_to_edges(nbr, cnt)


# --------------------------------------------------------------------------------------
# A2  node / edge features
# --------------------------------------------------------------------------------------


def compute_acceleration(next_pos, pos_seq):
    """utils.py:10-24: second difference of positions."""
    return next_pos - 2 * pos_seq[-1] + pos_seq[-2]


def get_nodes_vel(pos_seq, velocity_mean, velocity_std):
    """utils.py:27-40: normalised first differences, flattened per node oldest -> newest."""
    vel = pos_seq[1:] - pos_seq[:-1]                      # [k-1, N, 3]
    vel = (vel - velocity_mean) / velocity_std
    return vel.transpose(0, 1).reshape(vel.shape[1], -1)  # [N, 3(k-1)]


def get_edges_displacement(last_pos, senders, receivers, conn_r):
    """utils.py:43-61: (p_send - p_recv)/r and its 2-norm."""
    rel = (last_pos[senders] - last_pos[receivers]) / conn_r
    return torch.cat((rel, torch.norm(rel, dim=-1, keepdim=True)), dim=-1)


class GraphFeatures:
    """GraphBoundedMultimaterial(Control) (collate_utils.py:160-232) + the assembly in
    GraphAttributes._process_simple / process_collate (collate_utils.py:29-40, 68-87)."""

    def __init__(self, conn_r, stats, bounds, cartesian_idx, material_idx, control_idx=None,
                 max_neighbours=20, connectivity=get_connectivity):
        self.conn_r = conn_r
        self.stats = stats
        self.bounds = bounds
        self.cartesian_idx = list(cartesian_idx)
        self.material_idx = [material_idx] if isinstance(material_idx, int) else list(material_idx)
        self.control_idx = None if control_idx is None else list(control_idx)
        # NB collate_utils.py:34 drops max_neighbours on the noise-free path -> hard default 20.
        self.max_neighbours = max_neighbours
        self.connectivity = connectivity

    def compute_nodes(self, obs):
        pos_seq = obs[:, :, self.cartesian_idx]
        last = pos_seq[-1]
        vel = get_nodes_vel(pos_seq, self.stats["velocity_mean"], self.stats["velocity_std"])
        dist = torch.cat((last - self.bounds["lower_bounds"], self.bounds["upper_bounds"] - last), dim=1)
        feats = [vel, torch.clamp(dist / self.conn_r, min=-1, max=1), obs[-1][:, self.material_idx]]
        if self.control_idx is not None:
            ctrl = obs[-1][:, self.control_idx]
            feats.append((ctrl - self.stats["velocity_mean"]) / self.stats["velocity_std"])
        return torch.cat(feats, dim=-1)

    def compute_target(self, obs, tgt):
        """collate_utils.py:148-157 (unused inside rollouts, needed for training targets)."""
        acc = compute_acceleration(tgt, obs[:, :, self.cartesian_idx])
        return (acc - self.stats["acceleration_mean"]) / self.stats["acceleration_std"]

    def process(self, obs, tgt=None):
        last = obs[-1][:, self.cartesian_idx]
        nodes = self.compute_nodes(obs)
        senders, receivers = self.connectivity(last, self.conn_r, self.max_neighbours)
        edge_attr = get_edges_displacement(last, senders, receivers, self.conn_r)
        target = None if tgt is None else self.compute_target(obs, tgt)
        return nodes, edge_attr, senders, receivers, target

    def process_collate(self, batch):
        nodes_l, edges_l, index_l, tgt_l = [], [], [], []
        for g, (obs, tgt) in enumerate(batch):
            nodes, edge_attr, s, r, t = self.process(obs, tgt)
            nodes_l.append(nodes)
            edges_l.append(edge_attr)
            index_l.append(torch.stack((s, r)) + nodes.shape[0] * g)   # equal N per graph, as :76
            tgt_l.append(t)
        return torch.cat(nodes_l), torch.cat(edges_l), torch.cat(index_l, dim=1), torch.cat(tgt_l)


# --------------------------------------------------------------------------------------
# A3 / A4  graph-network blocks and the Encode-Process-Decode model (fp32 torch reference)
# --------------------------------------------------------------------------------------


def scatter_sum(src, idx, n_out):
    """gn_utils.py:60-64: zeros + scatter_add over dim 0 (sequential fp32 on CPU)."""
    out = torch.zeros(n_out, src.shape[1], dtype=src.dtype, device=src.device)
    return out.index_add_(0, idx, src)


class EdgeModel(nn.Module):
    """graph_networks.py:19-50: phi_edge(cat[edge_attr, nodes[recv], nodes[send]])."""

    def __init__(self, phi_edge, use_receiver_nodes=True, use_sender_nodes=True, use_context=False):
        super().__init__()
        self.use_receiver_nodes, self.use_sender_nodes, self.use_context = \
            use_receiver_nodes, use_sender_nodes, use_context
        self.phi_edge = phi_edge

    def forward(self, nodes, edge_attr, edge_index, context=None):
        if self.use_context and context is None:
            raise ValueError("EdgeModel use_globals set to True and globals not provided")
        parts = [edge_attr]
        if self.use_receiver_nodes:
            parts.append(nodes[edge_index[1]])
        if self.use_sender_nodes:
            parts.append(nodes[edge_index[0]])
        if self.use_context:
            parts.append(context.repeat(edge_attr.shape[0], 1))
        return self.phi_edge(torch.cat(parts, dim=-1))


class NodeModel(nn.Module):
    """graph_networks.py:53-85: phi_node(cat[sum_recv(edges), (sum_sent), (context), nodes])."""

    def __init__(self, phi_node, use_received_edges=True, use_sent_edges=False, use_context=False,
                 reduce="sum"):
        super().__init__()
        self.phi_node = phi_node
        self.use_received_edges, self.use_sent_edges, self.use_context = \
            use_received_edges, use_sent_edges, use_context
        self.reduce = reduce   # ignored by the reference aggregators (gn_utils.py:53-58): always sum

    def forward(self, nodes, edge_attr, edge_index, context=None):
        if self.use_context and context is None:
            raise ValueError("EdgeModel use_globals set to True and globals not provided")
        parts = []
        if self.use_received_edges:
            parts.append(scatter_sum(edge_attr, edge_index[1], nodes.shape[0]))
        if self.use_sent_edges:
            parts.append(scatter_sum(edge_attr, edge_index[0], nodes.shape[0]))
        if self.use_context:
            parts.append(context.repeat(nodes.shape[0], 1))
        parts.append(nodes)
        return self.phi_node(torch.cat(parts, dim=1))


class GraphNetwork(nn.Module):
    """graph_networks.py:88-97."""

    def __init__(self, phi_edge, phi_node, phi_context):
        super().__init__()
        self.phi_edge, self.phi_node, self.phi_global = phi_edge, phi_node, phi_context

    def forward(self, nodes, edge_attr, edge_index, context=None):
        raise NotImplementedError


class InteractionNetwork(GraphNetwork):
    """graph_networks.py:100-118. The same Sequential is registered twice (on self and inside
    edge_model/node_model) -> duplicate state-dict keys; node half sees the pre-residual edges."""

    def __init__(self, phi_edge, phi_node):
        super().__init__(phi_edge, phi_node, None)
        self.edge_model = EdgeModel(self.phi_edge)
        self.node_model = NodeModel(self.phi_node)

    def forward(self, nodes, edge_attr, edge_index):
        edge_out = self.edge_model(nodes, edge_attr, edge_index)
        return self.node_model(nodes, edge_out, edge_index), edge_out, edge_index


class GraphIndependent(GraphNetwork):
    """graph_networks.py:121-137."""

    def __init__(self, phi_edge=None, phi_node=None, phi_context=None):
        ident = lambda x: x  # noqa: E731
        super().__init__(phi_edge or ident, phi_node or ident, phi_context or ident)

    def forward(self, nodes, edge_attr, edge_index, context=None):
        return self.phi_node(nodes), self.phi_edge(edge_attr), self.phi_global(context)


def build_mlp(d_in, hidden, d_out, num_layers, norm):
    """epd_gnn.py:72-84: Linear,ReLU,(Linear,ReLU)x(num_layers-1),Linear,[LayerNorm]."""
    mods = [nn.Linear(d_in, hidden), nn.ReLU()]
    for _ in range(num_layers - 1):
        mods += [nn.Linear(hidden, hidden), nn.ReLU()]
    mods.append(nn.Linear(hidden, d_out))
    if norm:
        mods.append(nn.LayerNorm(d_out))
    return nn.Sequential(*mods)


class EncProcDecGNN(nn.Module):
    """epd_gnn.py:11-105.  Module construction order matches the reference so that
    torch.manual_seed(s) yields identical random-init weights."""

    def __init__(self, node_dim, edge_dim, out_dim, hidden_size, num_layers, m_steps,
                 norm_type="LayerNorm"):
        super().__init__()
        assert num_layers >= 2, "The number of layers num_layers must be at least 2"
        assert m_steps >= 1, "The number of m_steps message pasting steps must be at least 1"
        h = hidden_size
        self.encoder = GraphIndependent(phi_edge=build_mlp(edge_dim, h, h, num_layers, True),
                                        phi_node=build_mlp(node_dim, h, h, num_layers, True))
        self.processor = nn.ModuleList(
            InteractionNetwork(phi_edge=build_mlp(3 * h, h, h, num_layers, True),
                               phi_node=build_mlp(2 * h, h, h, num_layers, True))
            for _ in range(m_steps))
        self.decoder = build_mlp(h, h, out_dim, num_layers, False)

    def forward(self, nodes, edge_attr, edge_index):
        v, e, _ = self.encoder(nodes, edge_attr, edge_index)
        for block in self.processor:
            dv, de, _ = block(v, e, edge_index)
            v, e = dv + v, de + e          # residual on nodes and edges (epd_gnn.py:100-105)
        return self.decoder(v)


# --------------------------------------------------------------------------------------
# A5  integrator, rigid trajectory, rollout loop
# --------------------------------------------------------------------------------------


def get_position_from_prediction(stats, cartesian_idx, pred_acc, obs_seq):
    """rollout_utils.py:145-158: semi-implicit Euler, dt = 1 frame."""
    acc = pred_acc * stats["acceleration_std"] + stats["acceleration_mean"]
    last = obs_seq[-1][:, cartesian_idx]
    vel = (last - obs_seq[-2][:, cartesian_idx]) + acc
    return last + vel


def compute_particles_tmatrix(rotation, translation, ty_init, rigid_particles):
    """traj_utils.py:167-194 == rollout_utils.py:179-205: rigid particles rotated about X by
    ``rotation`` and shifted in Y by ``translation`` around ty_init, with y/z swapped and the
    offset taken as (ty_init - p).  Trig in f64 numpy, matrix entries cast to f32 (:171-175),
    the 4x4 @ 4xN product in f32."""
    c, s = np.cos(rotation), np.sin(rotation)
    T = torch.tensor([[1, 0, 0, ty_init[0]],
                      [0, c, -s, ty_init[1] + translation],
                      [0, s, c, ty_init[2]],
                      [0, 0, 0, 1]], dtype=torch.float32)
    q = torch.ones((4, rigid_particles.shape[0]))
    q[0] = ty_init[0] - rigid_particles[:, 0]
    q[1] = ty_init[1] - rigid_particles[:, 2]
    q[2] = ty_init[2] - rigid_particles[:, 1]
    w = T @ q
    return torch.stack((w[0], w[2], w[1]), dim=1)


def rigid_body_trajectory(traj_rot, traj_ty, horizon, ty_init, rigid_particles):
    """traj_utils.py:87-103 / rollout_utils.py:161-177 -> [horizon, N_rigid, 3]."""
    return torch.stack([compute_particles_tmatrix(traj_rot[i], traj_ty[i], ty_init, rigid_particles)
                        for i in range(horizon)])


class TrajectoryParam:
    """TrajectoryCMAsolver.set_sample_traj / interpolate_trajectory (traj_utils.py:199-228) with
    the ctor quirk scale_ty = (ty_init[0] - 0.3) / scale_rot (traj_utils.py:61)."""

    def __init__(self, rx_init_deg, ty_init, scale_rot, max_rot_deg, max_ty):
        self.rx_init = np.deg2rad(rx_init_deg)
        self.ty_init = ty_init
        self.scale_rot = scale_rot
        self.scale_ty = (ty_init[0] - 0.3) / scale_rot
        self.max_rot = np.deg2rad(max_rot_deg)
        self.max_ty = max_ty

    def x0_from_sample(self, sample_traj):
        d = sample_traj[2:] - sample_traj[1:-1]
        return np.concatenate((np.deg2rad(d[:, 0] / self.scale_rot), d[:, 1] / self.scale_ty))

    def interpolate(self, x):
        m = len(x) // 2
        inc_rot = np.clip(np.deg2rad(self.scale_rot * np.rad2deg(x[:m])), -self.max_rot, self.max_rot)
        inc_ty = np.clip(self.scale_ty * x[m:], -self.max_ty, self.max_ty)
        rot, ty = [self.rx_init], [0.0]
        for i in range(m):                 # sequential accumulation, as traj_utils.py:214-226
            rot.append(rot[-1] + inc_rot[i])
            ty.append(ty[-1] + inc_ty[i])
        return rot, ty


def pchip_slopes(xk, yk):
    """scipy.interpolate.PchipInterpolator._find_derivatives / _edge_case (scipy 1.18, _cubic.py; the reference
    calls scipy.interpolate.pchip_interpolate at traj_utils.py:413-414; scipy is unpinned in environment.yml)."""
    hk = xk[1:] - xk[:-1]
    mk = (yk[1:] - yk[:-1]) / hk
    n = len(yk)
    if n == 2:
        return np.array([mk[0], mk[0]])
    dk = np.zeros(n)
    for k in range(1, n - 1):
        m0, m1 = mk[k - 1], mk[k]
        if np.sign(m0) != np.sign(m1) or m0 == 0.0 or m1 == 0.0:
            dk[k] = 0.0
        else:
            w1 = 2 * hk[k] + hk[k - 1]
            w2 = hk[k] + 2 * hk[k - 1]
            dk[k] = 1.0 / ((w1 / m0 + w2 / m1) / (w1 + w2))

    def edge(h0, h1, m0, m1):
        d = ((2 * h0 + h1) * m0 - h0 * m1) / (h0 + h1)
        if np.sign(d) != np.sign(m0):
            return 0.0
        if np.sign(m0) != np.sign(m1) and abs(d) > 3.0 * abs(m0):
            return 3.0 * m0
        return d
    dk[0] = edge(hk[0], hk[1], mk[0], mk[1])
    dk[-1] = edge(hk[-1], hk[-2], mk[-1], mk[-2])
    return dk


def pchip_eval(xk, yk, xq):
    """CubicHermiteSpline coefficients (_cubic.py) and PPoly evaluation (_ppoly evaluate_poly1: the power basis,
    res = c3 + c2 s + c1 s^2 + c0 s^3 with s^k accumulated by repeated multiplication)."""
    xk = np.asarray(xk, dtype=np.float64)
    yk = np.asarray(yk, dtype=np.float64)
    dk = pchip_slopes(xk, yk)
    dx = xk[1:] - xk[:-1]
    slope = (yk[1:] - yk[:-1]) / dx
    t = (dk[:-1] + dk[1:] - 2 * slope) / dx
    c0, c1, c2, c3 = t / dx, (slope - dk[:-1]) / dx - t, dk[:-1], yk[:-1]
    out = np.empty(len(xq))
    for q, x in enumerate(xq):
        i = min(max(int(np.searchsorted(xk, x, side="right")) - 1, 0), len(xk) - 2)
        s = x - xk[i]
        res, z = 0.0, 1.0
        for c in (c3[i], c2[i], c1[i], c0[i]):
            res = res + c * z * 1.0
            z = z * s
        out[q] = res
    return out


class PchipParam:
    """InterpolatedCMAsolver (traj_utils.py:288-452): knots every ``nr_traj_points`` steps, parameters = scaled
    offsets of the knots from (rx_init, 0), curves by pchip interpolation; with the ctor quirk
    scale_ty = (ty_init[0] - 0.3) / scale_rot (traj_utils.py:61)."""

    def __init__(self, rx_init_deg, ty_init, scale_rot, max_rot_deg, max_ty, total_steps, traj_points):
        self.rx_init = np.deg2rad(rx_init_deg)
        self.ty_init = ty_init
        self.scale_rot = scale_rot
        self.scale_ty = (ty_init[0] - 0.3) / scale_rot
        self.max_rot = np.deg2rad(max_rot_deg)
        self.max_ty = max_ty
        self.horizon = total_steps
        self.nr = traj_points                       # steps between knots (self.nr_traj_points)
        self.n_knots = int(total_steps / traj_points)   # self.traj_points

    def x0_from_sample(self, sample_traj):
        """set_sample_traj (traj_utils.py:296-303)."""
        pts = sample_traj[list(range(self.nr, sample_traj.shape[0], self.nr))]
        return np.concatenate(((np.deg2rad(pts[:, 0]) - self.rx_init) / self.scale_rot,
                               (pts[:, 1] - self.ty_init[0]) / self.scale_ty))

    def interpolate(self, x):
        """interpolate_trajectory (traj_utils.py:393-416), type_interp='pchip'."""
        x = np.asarray(x, dtype=np.float64)
        m = len(x) // 2
        rot_pts = np.concatenate(([self.rx_init], self.rx_init + x[:m] * self.scale_rot))
        ty_pts = np.concatenate(([0.0], x[m:] * self.scale_ty))
        xk = np.arange(0, self.horizon + 1, self.nr).astype(np.float64)
        xq = np.arange(self.horizon).astype(np.float64)
        return pchip_eval(xk, rot_pts, xq), pchip_eval(xk, ty_pts, xq)

    def terms(self, x):
        """(vel_loss, acc_loss, interp_loss) of compute_loss (traj_utils.py:353-373, 418-452)."""
        rot, ty = self.interpolate(x)
        actions = np.stack((rot, ty), axis=1)
        vel = actions[1:] - actions[:-1]
        acc = actions[2:] - 2 * actions[1:-1] + actions[:-2]
        vel_loss = np.sqrt(np.sum((vel / np.array([1e-2, 4e-4])) ** 2))
        acc_loss = np.sqrt(np.sum((acc / np.array([2.2e-4, 1.45e-4])) ** 2))
        x = np.asarray(x, dtype=np.float64)
        r, t = x[:self.n_knots] * self.scale_rot, x[self.n_knots:] * self.scale_ty
        ineq = max((np.abs(r[1:] - r[:-1]) - self.max_rot * self.nr).max(),
                   (np.abs(t[1:] - t[:-1]) - self.max_ty * self.nr).max())
        return vel_loss, acc_loss, float(np.exp(ineq))

    def ineq_constraint(self, x):
        """traj_utils.py:375-391."""
        x = np.asarray(x, dtype=np.float64)
        actions = np.zeros((self.n_knots + 1, 2))
        actions[1:, 0] = x[:self.n_knots] * self.scale_rot
        actions[1:, 1] = x[self.n_knots:] * self.scale_ty
        vel = actions[1:] - actions[:-1]
        up = np.abs(vel) - np.array([self.max_rot * self.nr, self.max_ty * self.nr])
        return np.concatenate((up[:, 0] / self.scale_rot, up[:, 1] / self.scale_ty))


def increments_terms(param, x, rotation_limit=2.8973):
    """TrajectoryCMAsolver loss terms (traj_utils.py:230-246, 263-285): (vel_loss, acc_loss, bound penalty)."""
    rot, ty = param.interpolate(np.asarray(x, dtype=np.float64))
    actions = np.stack((np.array(rot), np.array(ty)), axis=1)
    vel = actions[1:] - actions[:-1]
    acc = actions[2:] - 2 * actions[1:-1] + actions[:-2]
    div = np.array([param.max_rot, param.max_ty])
    pen = 0.0
    if actions[:, 0].max() > param.rx_init + rotation_limit or actions[:, 0].min() < param.rx_init - rotation_limit:
        pen = 20.0
    return np.sqrt(np.sum((vel / div) ** 2)), np.sqrt(np.sum((acc / div) ** 2)), pen


# --------------------------------------------------------------------------------------
# N1  Sinkhorn divergence (objective tail of cma_objective) -- PARITY UNPINNED
# --------------------------------------------------------------------------------------


def sinkhorn_epsilon_schedule(diameter, blur, scaling, p=2):
    """geomloss sinkhorn_divergence.epsilon_schedule."""
    return ([diameter ** p] + [float(np.exp(e)) for e in
                               np.arange(p * np.log(diameter), p * np.log(blur), p * np.log(scaling))] + [blur ** p])


def sinkhorn_divergence(x, y, blur=0.05, scaling=0.5, dtype=np.float64):
    """geomloss.SamplesLoss(loss="sinkhorn", p=2, blur=.05)(x, y) for uniform weights, restated in numpy from the
    published tensorized algorithm (geomloss/sinkhorn_samples.py: softmin_tensorized, sinkhorn_tensorized;
    geomloss/sinkhorn_divergence.py: max_diameter, epsilon_schedule, sinkhorn_loop, sinkhorn_cost).  The reference
    calls it at traj_utils.py:276 / :439 (constructed at :69).  geomloss is an unpinned pip dependency
    (environment.yml:25) that is not installed here and not vendored in the reference: PARITY UNPINNED.  The one
    deliberate difference: the cost |x-y|^2/2 is formed from coordinate differences (geomloss expands it as
    |x|^2 - 2 x.y + |y|^2 in fp32, which loses about three digits for points 1e-2 apart at |x| ~ 0.5)."""
    from scipy.special import logsumexp
    x = np.asarray(x, dtype=dtype)
    y = np.asarray(y, dtype=dtype)
    n, m = len(x), len(y)

    def cost(a, b):
        d = a[:, None, :] - b[None, :, :]
        return 0.5 * (d * d).sum(-1)
    c_xy, c_yx, c_xx, c_yy = cost(x, y), cost(y, x), cost(x, x), cost(y, y)
    both = np.concatenate((x, y)).astype(np.float32)
    diameter = float(np.linalg.norm(both.max(0) - both.min(0)))
    eps_list = sinkhorn_epsilon_schedule(diameter, blur, scaling)
    a_log, b_log = np.full(n, -np.log(n), dtype=dtype), np.full(m, -np.log(m), dtype=dtype)

    def softmin(eps, c, h):
        return -eps * logsumexp(h[None, :] - c / eps, axis=1)
    eps = eps_list[0]
    g_ab, f_ba = softmin(eps, c_yx, a_log), softmin(eps, c_xy, b_log)
    f_aa, g_bb = softmin(eps, c_xx, a_log), softmin(eps, c_yy, b_log)
    for eps in eps_list:
        ft_ba, gt_ab = softmin(eps, c_xy, b_log + g_ab / eps), softmin(eps, c_yx, a_log + f_ba / eps)
        ft_aa, gt_bb = softmin(eps, c_xx, a_log + f_aa / eps), softmin(eps, c_yy, b_log + g_bb / eps)
        f_ba, g_ab = 0.5 * (f_ba + ft_ba), 0.5 * (g_ab + gt_ab)
        f_aa, g_bb = 0.5 * (f_aa + ft_aa), 0.5 * (g_bb + gt_bb)
    f_ba, g_ab = softmin(eps, c_xy, b_log + g_ab / eps), softmin(eps, c_yx, a_log + f_ba / eps)
    f_aa, g_bb = softmin(eps, c_xx, a_log + f_aa / eps), softmin(eps, c_yy, b_log + g_bb / eps)
    return float(np.mean(f_ba - f_aa) + np.mean(g_ab - g_bb))


def rollout(model, feats, state, trajectory, stats, horizon, record=None, teacher=None):
    """The loop of CMAESolver.cma_objective (traj_utils.py:119-152) == compute_rollout
    (rollout_utils.py:38-65) without the loss tail.  ``state`` = (obs [k,N,D], next_pos) is cloned.
    ``record(i, dict)`` receives per-step tensors (edges, features, acceleration, next_pos)."""
    obs = state[0].clone()
    cart, ctrl = feats.cartesian_idx, feats.control_idx
    rigid = obs[-1][:, feats.material_idx[0]] == 1
    with torch.no_grad():
        for i in range(horizon):
            if teacher is not None:
                obs = teacher(i).clone()
            new_rigid = obs[-1][rigid].clone()
            if i >= trajectory.shape[0]:   # reference quirk, traj_utils.py:128-129
                new_rigid[:, ctrl] = obs[-1][rigid][:, cart]
            else:
                new_rigid[:, ctrl] = trajectory[i] - obs[-1][rigid][:, cart]
            obs[-1][rigid] = new_rigid
            nodes, edge_attr, s, r, _ = feats.process(obs)
            pred = model(nodes, edge_attr, torch.stack((s, r)))
            nxt = get_position_from_prediction(stats, cart, pred, obs)
            if record is not None:
                record(i, dict(obs=obs, nodes=nodes, edge_attr=edge_attr, senders=s, receivers=r,
                               pred=pred, next_pos=nxt))
            obs[:-1] = obs[1:].clone()
            obs[-1][:, cart] = nxt
            if i < trajectory.shape[0]:
                new_rigid[:, cart] = trajectory[i]
            obs[-1][rigid] = new_rigid
    return obs
