"""Drop-in for gnn_manip/utils/utils.py (reference file:line cited per function), running on the B200.

Tensors may live on the CPU (as in the reference, whose graph/feature code is CPU-only) or on a CUDA device;
CPU inputs are copied to the current GPU, computed there and copied back -- there is no CPU implementation.
"""
import torch

from . import _lib
from ._lib import check, ptr, stream_ptr


class Graph:
    """Device-resident result of the radius-graph kernels (include/gnn_manip_b200.h, gm_radius_graph)."""

    def __init__(self, n_nodes, max_neighbours, device):
        i32 = dict(dtype=torch.int32, device=device)
        n, K = n_nodes, max_neighbours
        self.n_nodes, self.max_neighbours = n, K
        self.nbr = torch.empty((n, K), **i32)
        self.cnt = torch.empty((n,), **i32)
        self.edge_ptr = torch.empty((n + 1,), **i32)
        self.recv_ptr = torch.empty((n + 1,), **i32)
        self.esend = torch.empty((n * K,), **i32)
        self.erecv = torch.empty((n * K,), **i32)
        self.eref = torch.empty((n * K,), **i32)
        self.n_edges_dev = torch.zeros((1,), **i32)
        self._n_edges = None

    @property
    def n_edges(self):
        """Edge count (one device->host read, cached)."""
        if self._n_edges is None:
            self._n_edges = int(self.n_edges_dev.item())
        return self._n_edges

    def edge_index(self):
        """The reference's [2, E] int64 edge_index: row 0 senders, row 1 receivers, sender-major and
        distance-minor (gnn_manip/utils/utils.py:87-91, collate_utils.py:75-84)."""
        E = self.n_edges
        out = torch.empty((2, E), dtype=torch.int64, device=self.nbr.device)
        if E > 0:
            check(_lib.lib().gm_edge_index_i64(ptr(self.nbr), ptr(self.cnt), ptr(self.edge_ptr), self.n_nodes,
                                               self.max_neighbours, E, ptr(out), stream_ptr(out.device)),
                  "gm_edge_index_i64")
        return out


def _positions_f32(pos, dev):
    p = torch.as_tensor(pos)
    if p.dim() != 2 or p.shape[1] != 3:
        raise ValueError(f"Expected a [N, 3] position array, got shape {tuple(p.shape)}")
    return p.to(device=dev, dtype=torch.float32).contiguous()


def radius_graph(pos, conn_r, max_neighbours=20, n_graphs=1):
    """Build the capped radius graph of `pos` [n_graphs*N, 3] on the GPU; returns a device-resident Graph."""
    dev = _lib.compute_device(pos if torch.is_tensor(pos) else None)
    p = _positions_f32(pos, dev)
    n = p.shape[0]
    if n == 0 or n % n_graphs != 0:
        raise ValueError(f"radius_graph: {n} nodes cannot be split into {n_graphs} graphs")
    L = _lib.lib()
    with torch.cuda.device(dev):
        g = Graph(n, int(max_neighbours), dev)
        ws_bytes = L.gm_radius_graph_workspace_bytes(n // n_graphs, n_graphs, int(max_neighbours))
        ws = torch.empty((ws_bytes,), dtype=torch.uint8, device=dev)
        check(L.gm_radius_graph(ptr(p), 3, n // n_graphs, n_graphs, float(conn_r), int(max_neighbours), ptr(g.nbr),
                                ptr(g.cnt), ptr(g.edge_ptr), ptr(g.recv_ptr), ptr(g.esend), ptr(g.erecv), ptr(g.eref),
                                ptr(g.n_edges_dev), ptr(ws), ws_bytes, stream_ptr(dev)), "gm_radius_graph")
        g._keepalive = (p, ws)
    return g


def get_connectivity(pos_nodes, conn_r, max_neighbours=20):
    """gnn_manip/utils/utils.py:64-93 -> (senders, receivers) int64, on the device of `pos_nodes`."""
    out_dev = pos_nodes.device if torch.is_tensor(pos_nodes) else torch.device("cpu")
    g = radius_graph(pos_nodes, conn_r, max_neighbours)
    ei = g.edge_index().to(out_dev)
    return ei[0], ei[1]


def compute_acceleration(next_pos, pos_seq):
    """gnn_manip/utils/utils.py:10-24 (three tensor ops; not a kernel)."""
    return next_pos - 2 * pos_seq[-1] + pos_seq[-2]


def get_nodes_vel(pos_seq, velocity_mean, velocity_std):
    """gnn_manip/utils/utils.py:27-40: pos_seq [k, N, 3] -> normalised velocities [N, 3(k-1)].
    Runs the node-feature kernel with an all-position observation layout and returns its velocity block."""
    out_dev = pos_seq.device
    dev = _lib.compute_device(pos_seq)
    obs = pos_seq.to(device=dev, dtype=torch.float32).contiguous()
    k, n, d = obs.shape
    if d != 3:
        raise ValueError("get_nodes_vel expects [k, N, 3] positions")
    stats = {"velocity_mean": velocity_mean, "velocity_std": velocity_std}
    phys = _lib.make_phys(1.0, stats, None, [0, 1, 2], 0, None)
    F = _lib.load().gm_node_dim(k, 0)
    nodes = torch.empty((n, F), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        check(_lib.lib().gm_node_features(ptr(obs), k, n, 3, ctypes_ref(phys), ptr(nodes), stream_ptr(dev)),
              "gm_node_features")
    return nodes[:, : 3 * (k - 1)].contiguous().to(out_dev)


def ctypes_ref(struct):
    import ctypes
    return ctypes.byref(struct)


def get_edges_displacement(last_pos, senders, receivers, conn_r):
    """gnn_manip/utils/utils.py:43-61: [(p_send - p_recv)/r, ||.||] -> [E, 4]."""
    out_dev = last_pos.device
    dev = _lib.compute_device(last_pos)
    p = _positions_f32(last_pos, dev)
    s = torch.as_tensor(senders).to(device=dev, dtype=torch.int32).contiguous()
    r = torch.as_tensor(receivers).to(device=dev, dtype=torch.int32).contiguous()
    E = s.numel()
    out = torch.empty((E, 4), dtype=torch.float32, device=dev)
    if E > 0:
        with torch.cuda.device(dev):
            check(_lib.lib().gm_edge_features(ptr(p), 3, ptr(s), ptr(r), None, E, float(conn_r), ptr(out),
                                              stream_ptr(dev)), "gm_edge_features")
    return out.to(out_dev)


_noise_calls = [0]


def random_walk_noise(pos_seq, noise_std, seed=None, return_samples=False):
    """gnn_manip/utils/utils.py:96-115: position noise [k, N, 3] of a random walk in velocity whose last-step velocity
    noise has standard deviation `noise_std` (zero noise on the first frame).  Drawn on the device with Philox4x32-10
    (the reference samples torch.distributions.Normal on the CPU: same distribution, different stream).  `seed`
    defaults to torch.initial_seed() advanced by one per call."""
    import ctypes
    out_dev = pos_seq.device
    dev = _lib.compute_device(pos_seq)
    p = pos_seq.to(device=dev, dtype=torch.float32).contiguous()
    k, n, d = p.shape
    if d != 3:
        raise ValueError("random_walk_noise expects [k, N, 3] positions")
    if seed is None:
        seed = (torch.initial_seed() + _noise_calls[0]) & 0xFFFFFFFFFFFFFFFF
        _noise_calls[0] += 1
    phys = _lib.make_phys(1.0, None, None, [0, 1, 2], 0, None)
    zeros = torch.zeros_like(p)
    noisy = torch.empty_like(p)
    samples = torch.empty((k - 1, n, 3), dtype=torch.float32, device=dev) if return_samples else None
    with torch.cuda.device(dev):
        check(_lib.lib().gm_random_walk_noise(ptr(zeros), k, n, 3, ctypes.byref(phys), None, float(noise_std),
                                              ctypes.c_uint64(int(seed)), ptr(noisy), None, ptr(samples),
                                              stream_ptr(dev)), "gm_random_walk_noise")
    if return_samples:
        return noisy.to(out_dev), samples.to(out_dev)
    return noisy.to(out_dev)
