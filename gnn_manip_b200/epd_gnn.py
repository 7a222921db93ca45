"""Drop-in for gnn_manip/models/epd_gnn.py: EncProcDecGNN with the reference constructor, forward signature and
342-key state-dict layout (so dataset/trained_model.pth loads with load_state_dict(strict=True)); forward runs the
hand-written sm_100a kernels through the C ABI (gm_model_forward).  Inference only; no CPU implementation.
"""
import ctypes
import os

import torch
from torch import nn

from . import _lib
from ._lib import check, ptr, stream_ptr
from .graph_networks import GraphIndependent, InteractionNetwork, recv_csr


class PackedModel:
    """Device-side packed weights (gm_model) of one EncProcDecGNN on one device."""

    def __init__(self, module, dev):
        L = _lib.lib()
        dims = (module.node_dim, module.edge_dim, module.out_dim, module.hidden_size, module.num_layers, module.m_steps)
        flat = module.flat_parameters().to(device=dev, dtype=torch.float32).contiguous()
        n_params = L.gm_model_num_params(*dims)
        if flat.numel() != n_params:
            raise RuntimeError(f"EncProcDecGNN has {flat.numel()} unique parameters, the C ABI expects {n_params}")
        blob_bytes = L.gm_model_blob_bytes(*dims)
        self.blob = torch.empty((blob_bytes + 1024,), dtype=torch.uint8, device=dev)
        off = (-self.blob.data_ptr()) % 1024
        self.handle = ctypes.c_void_p()
        with torch.cuda.device(dev):
            check(L.gm_model_create(ctypes.byref(self.handle), ptr(flat), flat.numel(), *dims,
                                    ctypes.c_void_p(self.blob.data_ptr() + off), blob_bytes, stream_ptr(dev)),
                  "gm_model_create")
            torch.cuda.current_stream(dev).synchronize()   # `flat` may be freed after this
        self.device = dev
        self.supports_tc = bool(L.gm_model_supports_tc(self.handle))

    def __del__(self):
        try:
            if getattr(self, "handle", None):
                _lib.load().gm_model_destroy(self.handle)
                self.handle = None
        except Exception:
            pass


class EncProcDecGNN(nn.Module):
    """gnn_manip/models/epd_gnn.py:11-105.  Modules are constructed in the reference's order, so
    torch.manual_seed(s) followed by the constructor yields the reference's random initialisation."""

    def __init__(self, node_dim, edge_dim, out_dim, hidden_size, num_layers, m_steps, norm_type="LayerNorm"):
        super().__init__()
        assert num_layers >= 2, "The number of layers num_layers must be at least 2"
        assert m_steps >= 1, "The number of m_steps message pasting steps must be at least 1"
        if norm_type != "LayerNorm":
            raise NotImplementedError("only norm_type='LayerNorm' (the reference default) is implemented")
        self.node_dim, self.edge_dim, self.out_dim = node_dim, edge_dim, out_dim
        self.hidden_size, self.num_layers, self.m_steps = hidden_size, num_layers, m_steps
        h = hidden_size
        self.encoder = GraphIndependent(phi_edge=self._build_mlp(edge_dim, h, h, num_layers, True),
                                        phi_node=self._build_mlp(node_dim, h, h, num_layers, True))
        self.processor = nn.ModuleList()
        for _ in range(m_steps):
            self.processor.append(InteractionNetwork(phi_edge=self._build_mlp(3 * h, h, h, num_layers, True),
                                                     phi_node=self._build_mlp(2 * h, h, h, num_layers, True)))
        self.decoder = self._build_mlp(h, h, out_dim, num_layers, False)
        self.impl = int(os.environ.get("GNN_MANIP_B200_IMPL", _lib.GM_IMPL_AUTO))
        self.staged = False      # True: forward() calls gm_encode / gm_mp_step / gm_decode instead of gm_model_forward
        self._packed = {}      # device -> (version key, PackedModel)
        self._fwd_cache = {}   # device -> {"ws": (key, tensor), "csr": (key, (recv_ptr, esend, erecv, eref, n_edges))}

    @staticmethod
    def _build_mlp(in_dim, hidden, out_dim, num_layers, norm):
        """epd_gnn.py:72-84."""
        mods = [nn.Linear(in_dim, hidden), nn.ReLU()]
        for _ in range(num_layers - 1):
            mods += [nn.Linear(hidden, hidden), nn.ReLU()]
        mods.append(nn.Linear(hidden, out_dim))
        if norm:
            mods.append(nn.LayerNorm(out_dim))
        return nn.Sequential(*mods)

    # ---- packing -------------------------------------------------------------------------------------
    def _sequentials(self):
        yield self.encoder.phi_edge
        yield self.encoder.phi_node
        for blk in self.processor:
            yield blk.phi_edge
            yield blk.phi_node
        yield self.decoder

    def _ordered_parameters(self):
        """Unique parameters in the C ABI's flat order (include/gnn_manip_b200.h, gm_model_create)."""
        for seq in self._sequentials():
            norm = None
            for m in seq:
                if isinstance(m, nn.Linear):
                    yield m.weight
                    yield m.bias
                elif isinstance(m, nn.LayerNorm):
                    norm = m
            if norm is not None:
                yield norm.weight
                yield norm.bias

    def flat_parameters(self):
        return torch.cat([p.detach().reshape(-1).float() for p in self._ordered_parameters()])

    def packed(self, dev):
        key = tuple((p.data_ptr(), p._version) for p in self._ordered_parameters())
        hit = self._packed.get(dev)
        if hit is None or hit[0] != key:
            hit = (key, PackedModel(self, dev))
            self._packed[dev] = hit
        return hit[1]

    # ---- forward -------------------------------------------------------------------------------------
    @torch.no_grad()
    def forward(self, nodes, edge_attr, edge_index):
        """epd_gnn.py:86-98: nodes [N, node_dim], edge_attr [E, edge_dim], edge_index [2, E] (row 0 senders,
        row 1 receivers) -> normalised acceleration [N, out_dim] on the device of `nodes`."""
        out_dev = nodes.device
        dev = _lib.compute_device(nodes)
        if nodes.dim() != 2 or nodes.shape[1] != self.node_dim:
            raise RuntimeError(f"nodes must be [N, {self.node_dim}], got {tuple(nodes.shape)}")
        if edge_attr.dim() != 2 or edge_attr.shape[1] != self.edge_dim or edge_index.shape != (2, edge_attr.shape[0]):
            raise RuntimeError(f"edge_attr must be [E, {self.edge_dim}] and edge_index [2, E]; got "
                               f"{tuple(edge_attr.shape)} and {tuple(edge_index.shape)}")
        pm = self.packed(dev)
        L = _lib.lib()
        x = nodes.detach().to(device=dev, dtype=torch.float32).contiguous()
        ea = edge_attr.detach().to(device=dev, dtype=torch.float32).contiguous()
        N, E = x.shape[0], ea.shape[0]
        pred = torch.empty((N, self.out_dim), dtype=torch.float32, device=dev)
        cache = self._fwd_cache.setdefault(dev, {})
        with torch.cuda.device(dev):
            # a drop-in caller runs this once per rollout step: the workspace is kept per (N, E), the receiver CSR per
            # edge_index tensor (same storage and version -> same graph), the packed weights per parameter version
            # (the cache holds the edge_index tensor itself, so its storage cannot be recycled for another graph)
            hit = cache.get("csr")
            if hit is None or hit[0] is not edge_index or hit[1] != (edge_index._version, N):
                recv_ptr, esend, erecv, eref = recv_csr(edge_index[0], edge_index[1], N, dev)
                hit = (edge_index, (edge_index._version, N),
                       (recv_ptr, esend, erecv, eref, torch.tensor([E], dtype=torch.int32, device=dev)))
                cache["csr"] = hit
            recv_ptr, esend, erecv, eref, n_edges = hit[2]
            ws_bytes = L.gm_forward_workspace_bytes(pm.handle, N, E, self.impl)
            wkey = (N, E, self.impl, ws_bytes)
            hit = cache.get("ws")
            if hit is None or hit[0] != wkey:
                hit = (wkey, torch.empty((ws_bytes,), dtype=torch.uint8, device=dev))
                cache["ws"] = hit
            ws = hit[1]
            if E == 0:
                ea = torch.zeros((1, self.edge_dim), dtype=torch.float32, device=dev)
            if not self.staged:
                check(L.gm_model_forward(pm.handle, self.impl, ptr(x), ptr(ea), None, 0, 0.0, N, E, ptr(n_edges),
                                         ptr(recv_ptr), ptr(esend), ptr(erecv), ptr(eref), ptr(pred), ptr(ws), ws_bytes,
                                         stream_ptr(dev)), "gm_model_forward")
            else:
                # the reference's own structure (epd_gnn.py:88-96): encoder, one call per processor block, decoder
                st = stream_ptr(dev)
                check(L.gm_encode(pm.handle, self.impl, ptr(x), ptr(ea), None, 0, 0.0, N, E, ptr(n_edges), ptr(recv_ptr),
                                  ptr(esend), ptr(erecv), ptr(eref), ptr(ws), ws_bytes, st), "gm_encode")
                for step in range(self.m_steps):
                    check(L.gm_mp_step(pm.handle, self.impl, step, N, E, ptr(n_edges), ptr(recv_ptr), ptr(esend),
                                       ptr(erecv), ptr(ws), ws_bytes, st), "gm_mp_step")
                check(L.gm_decode(pm.handle, self.impl, N, E, ptr(pred), ptr(ws), ws_bytes, st), "gm_decode")
        return pred.to(out_dev)
