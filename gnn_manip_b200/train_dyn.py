"""Drop-in for the hot part of examples/train_dyn.py: `loss_batch` (:49-72) -- forward, L1Loss(sum)/N, backward, Adam --
for batches of one-step graphs, plus the data path that feeds it.

What changes underneath:
  * a batch is assembled ON THE DEVICE: the samples' observation windows are stacked along the particle axis, the
    random-walk noise (utils.py:96-115), node features, training targets and ONE batched radius graph (disjoint graphs,
    per-graph node offsets as collate_utils.py:75-76) are kernels; the reference builds each sample's graph on the CPU in
    Dataset.__getitem__ (coffee_dataset.py:107-113);
  * forward and backward are the tcgen05 kernels of csrc/train_kernels.cu driven by gm_train_forward_backward; autograd
    is not involved;
  * data parallelism: one process per GPU, every rank runs the same step on its own batch, the flat fp32 gradient
    (1 591 299 elements) is all-reduced over NCCL -- per bucket (decoder, processor blocks 9..0, encoders) on a side
    stream while the backward of the next block runs, or in one call -- averaged, and Adam is applied redundantly.
"""
import ctypes

import torch

from . import _lib
from ._lib import GRAD_READY_FN, check, gm_train_desc, ptr, stream_ptr


class Batch:
    """A device-resident training batch: B graphs of N particles as one disjoint graph."""

    def __init__(self, nodes, pos, graph, target, row_mask, n_graphs, conn_r):
        self.nodes, self.pos, self.graph, self.target, self.row_mask = nodes, pos, graph, target, row_mask
        self.n_graphs, self.conn_r = n_graphs, float(conn_r)


def make_batch(graph_attr, samples, device=None, seed=None, sand_only=False):
    """[(obs [k, N, D], next_pos [N, 3]), ...] -> Batch.  Follows GraphAttributes.process / process_collate
    (collate_utils.py:23-87) for equal-N samples: noise (if graph_attr.noise_std is set), node features, edges,
    normalised target acceleration.  sand_only: the `use_updated_loss` row selection (train_dyn.py:57-61)."""
    from .utils import radius_graph
    L = _lib.lib()
    dev = torch.device(device) if device is not None else _lib.compute_device(samples[0][0])
    f32 = dict(device=dev, dtype=torch.float32)
    B = len(samples)
    obs = torch.cat([s[0].to(**f32) for s in samples], dim=1).contiguous()          # [k, B*N, D]
    tgt = torch.cat([s[1].to(**f32) for s in samples], dim=0).contiguous()          # [B*N, 3]
    k, n, D = obs.shape
    phys = graph_attr.phys()
    cart = list(graph_attr.cartesian_idx)
    with torch.cuda.device(dev):
        if graph_attr.noise_std is not None:
            if seed is None:
                seed = int(torch.randint(0, 2 ** 62, (1,)).item())
            noisy_obs, noisy_tgt = torch.empty_like(obs), torch.empty_like(tgt)
            check(L.gm_random_walk_noise(ptr(obs), k, n, D, ctypes.byref(phys), ptr(tgt), float(graph_attr.noise_std),
                                         ctypes.c_uint64(seed), ptr(noisy_obs), ptr(noisy_tgt), None, stream_ptr(dev)),
                  "gm_random_walk_noise")
            obs, tgt = noisy_obs, noisy_tgt
            K = graph_attr._noisy_max_neighbours()
        else:
            K = graph_attr._simple_max_neighbours
        F = L.gm_node_dim(k, phys.use_control)
        nodes = torch.empty((n, F), **f32)
        check(L.gm_node_features(ptr(obs), k, n, D, ctypes.byref(phys), ptr(nodes), stream_ptr(dev)), "gm_node_features")
        target = torch.empty((n, 3), **f32)
        check(L.gm_training_target(ptr(obs), k, n, D, ctypes.byref(phys), ptr(tgt), ptr(target), stream_ptr(dev)),
              "gm_training_target")
        pos = obs[-1][:, cart].contiguous()
        graph = radius_graph(pos, graph_attr.conn_r, K, n_graphs=B)
        mat = graph_attr.material_idx[0] if isinstance(graph_attr.material_idx, (list, tuple)) else graph_attr.material_idx
        row_mask = (obs[-1][:, mat] < 0.5).to(torch.uint8).contiguous() if sand_only else None
    return Batch(nodes, pos, graph, target, row_mask, B, graph_attr.conn_r)


def average_gradients(grads, buckets=None):
    """Data-parallel gradient exchange on a flat gradient vector: sum over the ranks (one all-reduce, or one per
    (offset, count) bucket), divided by the world size.  No-op without torch.distributed.  Works on any backend
    (NCCL on the GPUs; the gloo tests run it on CPU tensors)."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return grads
    if buckets is None:
        dist.all_reduce(grads)
    else:
        works = [dist.all_reduce(grads[o: o + c], async_op=True) for o, c in buckets]
        for w in works:
            w.wait()
    grads.div_(dist.get_world_size())
    return grads


class Trainer:
    """Optimiser state + training plan of one EncProcDecGNN on one device.

    trainer = Trainer(model, max_nodes, max_edges, lr=1e-4)          # examples/train_dyn.py:98: Adam(lr)
    loss = trainer.loss_batch(batch)                                  # forward, loss, backward, all-reduce, Adam
    trainer.export_to(model)                                          # write the parameters back into the nn.Module
    """

    def __init__(self, model, max_nodes, max_edges, lr=1e-4, betas=(0.9, 0.999), eps=1e-8, device=None,
                 bucketed_allreduce=True):
        L = _lib.lib()
        dev = torch.device(device) if device is not None else _lib.compute_device(None)
        self.device, self.model = dev, model
        self.lr, self.betas, self.eps = float(lr), (float(betas[0]), float(betas[1])), float(eps)
        self.step_count = 0
        self.params = model.flat_parameters().to(device=dev, dtype=torch.float32).contiguous()
        self.grads = torch.zeros_like(self.params)
        self.exp_avg, self.exp_avg_sq = torch.zeros_like(self.params), torch.zeros_like(self.params)
        d = gm_train_desc()
        d.node_dim, d.edge_dim, d.out_dim = model.node_dim, model.edge_dim, model.out_dim
        d.hidden, d.num_layers, d.m_steps = model.hidden_size, model.num_layers, model.m_steps
        d.max_nodes, d.max_edges = int(max_nodes), int(max_edges)
        d.params, d.grads = self.params.data_ptr(), self.grads.data_ptr()
        self.desc = d
        ws_bytes = L.gm_train_workspace_bytes(ctypes.byref(d))
        if ws_bytes == 0:
            raise _lib.GmError("the training kernels cover hidden_size=128, num_layers=2 (the reference configuration)")
        self.workspace = torch.empty((ws_bytes + 1024,), dtype=torch.uint8, device=dev)
        off = (-self.workspace.data_ptr()) % 1024
        self.handle = ctypes.c_void_p()
        with torch.cuda.device(dev):
            check(L.gm_train_create(ctypes.byref(self.handle), ctypes.byref(d),
                                    ctypes.c_void_p(self.workspace.data_ptr() + off), ws_bytes, stream_ptr(dev)),
                  "gm_train_create")
        self.loss = torch.zeros((1,), dtype=torch.float32, device=dev)
        self.buckets = []
        off_c, cnt_c = ctypes.c_size_t(), ctypes.c_size_t()
        for b in range(L.gm_train_num_buckets(self.handle)):
            check(L.gm_train_bucket(self.handle, b, ctypes.byref(off_c), ctypes.byref(cnt_c)), "gm_train_bucket")
            self.buckets.append((off_c.value, cnt_c.value))
        self.bucketed_allreduce = bool(bucketed_allreduce)
        self._comm_stream = None
        self._pending = []

    # ---- collectives -----------------------------------------------------------------------------------
    @staticmethod
    def _dist():
        import torch.distributed as dist
        return dist if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1 else None

    def _bucket_ready(self, user, bucket, offset, count):
        """Host callback fired by gm_train_forward_backward once bucket `bucket` of the gradient is enqueued: order an
        all-reduce of that slice behind it on the communication stream (overlaps the backward of the next block)."""
        dist = self._dist()
        compute = torch.cuda.current_stream(self.device)
        ev = torch.cuda.Event()
        ev.record(compute)
        self._comm_stream.wait_event(ev)
        with torch.cuda.stream(self._comm_stream):
            self._pending.append(dist.all_reduce(self.grads[offset: offset + count], async_op=True))

    def all_reduce_gradients(self):
        """One all-reduce over the whole flat gradient (used when the per-bucket overlap is off)."""
        average_gradients(self.grads)

    # ---- the step --------------------------------------------------------------------------------------
    def forward_backward(self, batch, backward=True):
        """Loss (device scalar tensor) of `batch`; with backward=True the gradients of every parameter land in
        self.grads (already all-reduced and averaged under torch.distributed)."""
        L = _lib.lib()
        g = batch.graph
        dist = self._dist() if backward else None
        cb = None
        if dist is not None and self.bucketed_allreduce:
            if self._comm_stream is None:
                self._comm_stream = torch.cuda.Stream(self.device)
            self._pending = []
            cb = GRAD_READY_FN(self._bucket_ready)
        n, max_edges = batch.nodes.shape[0], g.esend.numel()
        with torch.cuda.device(self.device):
            check(L.gm_train_forward_backward(self.handle, ptr(batch.nodes), ptr(batch.pos), batch.pos.shape[1],
                                              float(self._conn_r(batch)), n, max_edges, ptr(g.n_edges_dev),
                                              ptr(g.recv_ptr), ptr(g.esend), ptr(g.erecv), ptr(g.eref), ptr(g.edge_ptr),
                                              ptr(batch.target), ptr(batch.row_mask), ptr(self.loss),
                                              1 if backward else 0, cb, None, stream_ptr(self.device)),
                  "gm_train_forward_backward")
            if dist is not None:
                if cb is not None:
                    for w in self._pending:
                        w.wait()                       # orders the compute stream behind the collectives
                    torch.cuda.current_stream(self.device).wait_stream(self._comm_stream)
                    self.grads.div_(dist.get_world_size())
                else:
                    self.all_reduce_gradients()
        return self.loss

    @staticmethod
    def _conn_r(batch):
        return batch.conn_r

    def optimizer_step(self):
        self.step_count += 1
        L = _lib.lib()
        with torch.cuda.device(self.device):
            check(L.gm_adam_step(ptr(self.params), ptr(self.grads), ptr(self.exp_avg), ptr(self.exp_avg_sq),
                                 self.params.numel(), self.lr, self.betas[0], self.betas[1], self.eps, self.step_count,
                                 stream_ptr(self.device)), "gm_adam_step")
            check(L.gm_train_sync_weights(self.handle, stream_ptr(self.device)), "gm_train_sync_weights")

    def loss_batch(self, batch, test=False):
        """examples/train_dyn.py:49-72: returns the loss tensor (no host sync); trains unless test=True."""
        loss = self.forward_backward(batch, backward=not test)
        if not test:
            self.optimizer_step()
        return loss

    def prediction(self, n_rows):
        """Normalised accelerations [n_rows, out_dim] of the last forward (a copy)."""
        L = _lib.lib()
        addr = L.gm_train_prediction(self.handle)
        base = self.workspace.data_ptr()
        nbytes = n_rows * self.model.out_dim * 4
        return self.workspace[addr - base: addr - base + nbytes].view(torch.float32).reshape(n_rows, -1).clone()

    def activation(self, which, step, n_rows):
        """Test aid: a copy of a saved post-ReLU activation [n_rows, 128] (gm_train_activation)."""
        addr = _lib.lib().gm_train_activation(self.handle, int(which), int(step))
        base = self.workspace.data_ptr()
        return self.workspace[addr - base: addr - base + n_rows * 128 * 4].view(torch.float32).reshape(n_rows, 128).clone()

    def export_to(self, model=None):
        """Write the flat parameters back into the nn.Module (state-dict layout unchanged: torch.save works as in
        train_dyn.py:20-22)."""
        model = model or self.model
        flat = self.params.detach().cpu()
        off = 0
        with torch.no_grad():
            for p in model._ordered_parameters():
                n = p.numel()
                p.copy_(flat[off: off + n].reshape(p.shape))
                off += n
        return model

    def load_from(self, model=None):
        model = model or self.model
        self.params.copy_(model.flat_parameters().to(self.device))
        with torch.cuda.device(self.device):
            check(_lib.lib().gm_train_sync_weights(self.handle, stream_ptr(self.device)), "gm_train_sync_weights")

    def close(self):
        if getattr(self, "handle", None):
            _lib.load().gm_train_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
