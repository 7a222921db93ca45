"""Drop-in for gnn_manip/utils/collate_utils.py: the GraphAttributes family with the reference constructors and
process / process_collate / evaluate signatures.  Graph build and feature build run as CUDA kernels (rows A1, A2 of
SURVEY.md section 8); results come back on the device of the observation tensor (CPU in the reference's callers).
"""
import ctypes

import torch

from . import _lib
from ._lib import check, ptr, stream_ptr
from .utils import (compute_acceleration, get_edges_displacement, get_nodes_vel, radius_graph,
                    random_walk_noise)


class GraphAttributes:
    """collate_utils.py:13-124."""

    def __init__(self, conn_r, stats, cartesian_idx, action_idx, bounds, noise_std):
        self.conn_r = conn_r
        self.stats = stats
        self.cartesian_idx = cartesian_idx
        self.action_idx = action_idx
        self.bounds = bounds
        self.noise_std = noise_std

    # NB collate_utils.py:34 drops max_neighbours on the noise-free path: the cap is utils.py:64's default 20.
    _simple_max_neighbours = 20

    def process(self, obs, tgt):
        if self.noise_std is None:
            return self._process_simple(obs, tgt)
        return self._process_noisy(obs, tgt)

    def _process_simple(self, obs, tgt):
        out_dev = obs.device
        last_pos = obs[-1][:, self.cartesian_idx]
        nodes = self.compute_nodes(obs)
        graph = radius_graph(last_pos, self.conn_r, self._simple_max_neighbours)
        ei = graph.edge_index()
        senders, receivers = ei[0].to(out_dev), ei[1].to(out_dev)
        edge_attr = self.compute_edges(obs, senders, receivers)
        nodes_tgt = self.compute_target(obs, tgt)
        return nodes, edge_attr, senders, receivers, nodes_tgt

    def _noisy(self, obs, tgt):
        """collate_utils.py:41-56 / 169-183: (noisy_obs, noisy_tgt) -- random-walk noise on the cartesian columns, its last
        frame added to the target."""
        pos_seq = obs[:, :, self.cartesian_idx]
        noise = random_walk_noise(pos_seq, self.noise_std).to(obs.device)
        noisy_obs = obs.clone()
        noisy_obs[:, :, self.cartesian_idx] = obs[:, :, self.cartesian_idx] + noise
        return noisy_obs, tgt + noise[-1]

    def _process_noisy(self, obs, tgt):
        """collate_utils.py:41-66 (the base class drops max_neighbours here too: cap 20)."""
        noisy_obs, noisy_tgt = self._noisy(obs, tgt)
        out_dev = obs.device
        last_pos = noisy_obs[-1][:, self.cartesian_idx]
        nodes = self.compute_nodes(noisy_obs)
        ei = radius_graph(last_pos, self.conn_r, self._noisy_max_neighbours()).edge_index()
        senders, receivers = ei[0].to(out_dev), ei[1].to(out_dev)
        edge_attr = self.compute_edges(noisy_obs, senders, receivers)
        return nodes, edge_attr, senders, receivers, self.compute_target(noisy_obs, noisy_tgt)

    def _noisy_max_neighbours(self):
        return self._simple_max_neighbours

    def process_collate(self, batch):
        """collate_utils.py:68-87: concatenates per-sample graphs with node offset N*i (equal N per sample)."""
        nodes_l, edges_l, index_l, tgt_l = [], [], [], []
        for i, sample in enumerate(batch):
            nodes, edge_attr, senders, receivers, nodes_tgt = self.process(sample[0], sample[1])
            index_l.append(torch.stack((senders, receivers)).long() + nodes.shape[0] * i)
            nodes_l.append(nodes)
            edges_l.append(edge_attr)
            tgt_l.append(nodes_tgt)
        return torch.cat(nodes_l), torch.cat(edges_l), torch.cat(index_l, dim=1), torch.cat(tgt_l)

    def evaluate(self, batch):
        """collate_utils.py:89-112."""
        nodes, edge_attr, edge_index, nodes_tgt = self.process_collate(batch)
        pos_seq = torch.stack([b[0] for b in batch])
        tgt_next_pos = torch.stack([b[1] for b in batch])
        return nodes, edge_attr, edge_index, nodes_tgt, pos_seq, tgt_next_pos

    def compute_nodes(self, obs):
        raise NotImplementedError

    def compute_edges(self, obs, senders, receivers):
        raise NotImplementedError

    def compute_target(self, obs, tgt):
        raise NotImplementedError


class GraphSimple(GraphAttributes):
    """collate_utils.py:127-157: nodes = last_pos || velocities."""

    def __init__(self, conn_r, stats, cartesian_idx, action_idx=None, bounds=None, noise=None):
        super().__init__(conn_r, stats, cartesian_idx, action_idx, bounds, noise)

    def compute_nodes(self, obs):
        pos_seq = obs[:, :, self.cartesian_idx]
        vel = get_nodes_vel(pos_seq, self.stats["velocity_mean"], self.stats["velocity_std"])
        return torch.cat((pos_seq[-1], vel), dim=-1)

    def compute_edges(self, obs, senders, receivers):
        return get_edges_displacement(obs[-1][:, self.cartesian_idx], senders, receivers, self.conn_r)

    def compute_target(self, obs, tgt):
        acc = compute_acceleration(tgt, obs[:, :, self.cartesian_idx])
        return (acc - self.stats["acceleration_mean"].to(acc.device)) / self.stats["acceleration_std"].to(acc.device)


class GraphBoundedMultimaterial(GraphSimple):
    """collate_utils.py:160-208: nodes = velocities || clamped boundary distances || material."""

    def __init__(self, conn_r, stats, cartesian_idx, material_idx, bounds, noise=None, max_neighbours=20):
        self.material_idx = material_idx
        self.max_neighbours = max_neighbours
        super().__init__(conn_r, stats, cartesian_idx, None, bounds, noise)

    _control_idx = None

    def _noisy_max_neighbours(self):
        return self.max_neighbours           # collate_utils.py:187: the noisy path of this class does pass the cap

    def phys(self):
        return _lib.make_phys(self.conn_r, self.stats, self.bounds, self.cartesian_idx, self.material_idx,
                              self._control_idx)

    def compute_nodes(self, obs):
        out_dev = obs.device
        dev = _lib.compute_device(obs)
        o = obs.to(device=dev, dtype=torch.float32).contiguous()
        k, n, D = o.shape
        phys = self.phys()
        F = _lib.load().gm_node_dim(k, phys.use_control)
        nodes = torch.empty((n, F), dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            check(_lib.lib().gm_node_features(ptr(o), k, n, D, ctypes.byref(phys), ptr(nodes), stream_ptr(dev)),
                  "gm_node_features")
        return nodes.to(out_dev)


class GraphBoundedMultimaterialControl(GraphBoundedMultimaterial):
    """collate_utils.py:211-232: + normalised control columns."""

    def __init__(self, conn_r, stats, cartesian_idx, material_idx, control_idx, bounds, noise=None,
                 max_neighbours=20):
        self.control_idx = control_idx
        super().__init__(conn_r, stats, cartesian_idx, material_idx, bounds, noise, max_neighbours=max_neighbours)

    @property
    def _control_idx(self):
        return self.control_idx
