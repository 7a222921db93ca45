"""Drop-in for deps/torch-graphnet/torch_graphnet/graph_networks.py (+ utils/gn_utils.py): same class names,
constructor arguments, forward signatures and state-dict keys; the forward passes run hand-written CUDA kernels
through the C ABI (gm_mlp_forward: gather + concat + MLP + LayerNorm fused; gm_segment_sum: deterministic
receiver-sorted scatter_add).  Inference only (no autograd graph is recorded); there is no CPU implementation.
"""
import ctypes

import torch
from torch import nn

from . import _lib
from ._lib import check, ptr, stream_ptr


def _parse_phi(phi):
    """nn.Sequential(Linear, ReLU, ..., Linear[, LayerNorm]) -> (linears, layernorm | None)."""
    if not isinstance(phi, nn.Sequential):
        raise NotImplementedError("gnn_manip_b200 blocks need phi to be an nn.Sequential of Linear/ReLU[/LayerNorm]")
    mods = list(phi)
    ln = None
    if mods and isinstance(mods[-1], nn.LayerNorm):
        ln = mods.pop()
        if abs(ln.eps - 1e-5) > 0 or not ln.elementwise_affine:
            raise NotImplementedError("LayerNorm must be affine with eps=1e-5 (the reference configuration)")
    linears = []
    for i, m in enumerate(mods):
        if i % 2 == 0:
            if not isinstance(m, nn.Linear):
                raise NotImplementedError(f"phi[{i}] must be nn.Linear, got {type(m).__name__}")
            linears.append(m)
        elif not isinstance(m, nn.ReLU):
            raise NotImplementedError(f"phi[{i}] must be nn.ReLU, got {type(m).__name__}")
    if len(mods) % 2 == 0 or not linears:
        raise NotImplementedError("phi must end with a Linear (optionally followed by LayerNorm)")
    return linears, ln


def mlp_forward(phi, sources, indices, n_rows):
    """out[m] = phi(cat_s sources[s][indices[s][m] if indices[s] is not None else m]) on the GPU."""
    linears, ln = _parse_phi(phi)
    out_dev = sources[0].device
    dev = _lib.compute_device(sources[0])
    f32 = dict(device=dev, dtype=torch.float32)
    srcs = [s.detach().to(**f32).contiguous() for s in sources]
    idxs = [None if i is None else i.detach().to(device=dev, dtype=torch.int32).contiguous() for i in indices]
    # transposed weights are cached on the Sequential per (device, parameter versions)
    wkey = (str(dev),) + tuple((l.weight.data_ptr(), l.weight._version) for l in linears)
    hit = getattr(phi, "_gm_wt_cache", None)
    if hit is None or hit[0] != wkey:
        hit = (wkey, [l.weight.detach().to(**f32).t().contiguous() for l in linears])
        phi._gm_wt_cache = hit
    wts = hit[1]
    bs = [None if l.bias is None else l.bias.detach().to(**f32).contiguous() for l in linears]
    dims = [linears[0].in_features] + [l.out_features for l in linears]
    g = None if ln is None else ln.weight.detach().to(**f32).contiguous()
    b = None if ln is None else ln.bias.detach().to(**f32).contiguous()
    out = torch.empty((n_rows, dims[-1]), **f32)
    if n_rows == 0:
        return out.to(out_dev)
    nl, ns = len(linears), len(srcs)
    arr = lambda ts: (ctypes.c_void_p * len(ts))(*[0 if t is None else t.data_ptr() for t in ts])  # noqa: E731
    with torch.cuda.device(dev):
        check(_lib.lib().gm_mlp_forward(nl, arr(wts), arr(bs), (ctypes.c_int32 * (nl + 1))(*dims), ptr(g), ptr(b), ns,
                                        arr(srcs), arr(idxs), (ctypes.c_int32 * ns)(*[s.shape[1] for s in srcs]),
                                        n_rows, ptr(out), stream_ptr(dev)), "gm_mlp_forward")
    return out.to(out_dev)


def recv_csr(senders, receivers, n_nodes, dev):
    """Stable sort of an explicit edge list by receiver -> (recv_ptr, esend, erecv, eref) int32 on `dev`."""
    s = senders.detach().to(device=dev, dtype=torch.int32).contiguous()
    r = receivers.detach().to(device=dev, dtype=torch.int32).contiguous()
    E = s.numel()
    i32 = dict(dtype=torch.int32, device=dev)
    recv_ptr = torch.empty((n_nodes + 1,), **i32)
    esend, erecv, eref = (torch.empty((max(E, 1),), **i32) for _ in range(3))
    L = _lib.lib()
    ws_bytes = L.gm_recv_csr_workspace_bytes(n_nodes, E)
    ws = torch.empty((ws_bytes,), dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        check(L.gm_recv_csr(ptr(s), ptr(r), E, n_nodes, ptr(recv_ptr), ptr(esend), ptr(erecv), ptr(eref), ptr(ws),
                            ws_bytes, stream_ptr(dev)), "gm_recv_csr")
    return recv_ptr, esend, erecv, eref


def scatter_sum(edge_attr, index, n_nodes):
    """utils/gn_utils.py:60-64 (zeros + scatter_add over dim 0), as a receiver-sorted segmented sum that visits
    each node's edges in ascending edge id -- bit-identical to the CPU reference's sequential accumulation."""
    out_dev = edge_attr.device
    dev = _lib.compute_device(edge_attr)
    src = edge_attr.detach().to(device=dev, dtype=torch.float32).contiguous()
    arange = torch.arange(index.numel(), device=dev, dtype=torch.int32)
    recv_ptr, _, _, eref = recv_csr(arange, index, n_nodes, dev)
    agg = torch.empty((n_nodes, src.shape[1]), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        check(_lib.lib().gm_segment_sum(ptr(src), src.shape[1], ptr(recv_ptr), ptr(eref), n_nodes, ptr(agg),
                                        stream_ptr(dev)), "gm_segment_sum")
    return agg.to(out_dev)


def received_edges_to_node_aggregator(nodes, edge_attr, receivers, reduce="sum"):
    """utils/gn_utils.py:53-55 (the reference ignores `reduce`: always a sum)."""
    return scatter_sum(edge_attr, receivers, nodes.shape[0])


def sent_edges_to_node_aggregator(nodes, edge_attr, senders, reduce="sum"):
    """utils/gn_utils.py:56-58."""
    return scatter_sum(edge_attr, senders, nodes.shape[0])


class EdgeModel(nn.Module):
    """graph_networks.py:19-50: phi_edge(cat[edge_attr, nodes[receivers], nodes[senders], (context)])."""

    def __init__(self, phi_edge, use_receiver_nodes=True, use_sender_nodes=True, use_context=False):
        super().__init__()
        self._use_receiver_nodes = use_receiver_nodes
        self._use_sender_nodes = use_sender_nodes
        self._use_context = use_context
        self.phi_edge = phi_edge

    @torch.no_grad()
    def forward(self, nodes, edge_attr, edge_index, context=None):
        if self._use_context and context is None:
            raise ValueError("EdgeModel use_globals set to True and globals not provided")
        senders, receivers = edge_index[0], edge_index[1]
        srcs, idxs = [edge_attr], [None]
        if self._use_receiver_nodes:
            srcs.append(nodes)
            idxs.append(receivers)
        if self._use_sender_nodes:
            srcs.append(nodes)
            idxs.append(senders)
        if self._use_context:
            srcs.append(context.reshape(1, -1))
            idxs.append(torch.zeros(edge_attr.shape[0], dtype=torch.int32))
        return mlp_forward(self.phi_edge, srcs, idxs, edge_attr.shape[0])


class NodeModel(nn.Module):
    """graph_networks.py:53-85: phi_node(cat[sum_received(edges), (sum_sent(edges)), (context), nodes])."""

    def __init__(self, phi_node, use_received_edges=True, use_sent_edges=False, use_context=False, reduce="sum"):
        super().__init__()
        self.phi_node = phi_node
        self._use_received_edges = use_received_edges
        self._use_sent_edges = use_sent_edges
        self._use_context = use_context
        self._reduce = reduce

    @torch.no_grad()
    def forward(self, nodes, edge_attr, edge_index, context=None):
        if self._use_context and context is None:
            raise ValueError("EdgeModel use_globals set to True and globals not provided")
        senders, receivers = edge_index[0], edge_index[1]
        srcs, idxs = [], []
        if self._use_received_edges:
            srcs.append(received_edges_to_node_aggregator(nodes, edge_attr, receivers, self._reduce))
            idxs.append(None)
        if self._use_sent_edges:
            srcs.append(sent_edges_to_node_aggregator(nodes, edge_attr, senders, self._reduce))
            idxs.append(None)
        if self._use_context:
            srcs.append(context.reshape(1, -1))
            idxs.append(torch.zeros(nodes.shape[0], dtype=torch.int32))
        srcs.append(nodes)
        idxs.append(None)
        return mlp_forward(self.phi_node, srcs, idxs, nodes.shape[0])


class GraphNetwork(nn.Module):
    """graph_networks.py:88-97."""

    def __init__(self, phi_edge, phi_node, phi_context):
        super().__init__()
        self.phi_edge = phi_edge
        self.phi_node = phi_node
        self.phi_global = phi_context

    def forward(self, nodes, edge_attr, edge_index, context=None):
        raise NotImplementedError


class InteractionNetwork(GraphNetwork):
    """graph_networks.py:100-118.  phi_edge / phi_node are registered on the network *and* inside its
    EdgeModel / NodeModel, which is what gives the reference its duplicated state-dict keys
    (processor.i.phi_edge.* == processor.i.edge_model.phi_edge.*)."""

    def __init__(self, phi_edge, phi_node):
        super().__init__(phi_edge, phi_node, None)
        self.edge_model = EdgeModel(self.phi_edge)
        self.node_model = NodeModel(self.phi_node)

    @torch.no_grad()
    def forward(self, nodes, edge_attr, edge_index):
        edge_out = self.edge_model(nodes, edge_attr, edge_index)
        node_out = self.node_model(nodes, edge_out, edge_index)     # pre-residual edges (graph_networks.py:115-116)
        return node_out, edge_out, edge_index


class GraphIndependent(GraphNetwork):
    """graph_networks.py:121-137."""

    def __init__(self, phi_edge=None, phi_node=None, phi_context=None):
        super().__init__(phi_edge if phi_edge is not None else nn.Identity(),
                         phi_node if phi_node is not None else nn.Identity(),
                         phi_context if phi_context is not None else nn.Identity())

    @torch.no_grad()
    def forward(self, nodes, edge_attr, edge_index, context=None):
        def run(phi, x):
            if x is None or isinstance(phi, nn.Identity):
                return x
            return mlp_forward(phi, [x], [None], x.shape[0])
        return run(self.phi_node, nodes), run(self.phi_edge, edge_attr), run(self.phi_global, context)
