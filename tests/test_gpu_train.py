"""GPU suite (-m gpu) for rows E(2) / N3: the training step (forward with saved activations, L1 loss, backward, Adam) and
the device-side training data path, against torch autograd over the CPU oracle model (the same torch modules the
reference trains through) on the same seeded batch."""
import numpy as np
import pytest
import torch

from conftest import rel_l2
from oracle import scene
from test_gpu_parity import oracle_features, seeded_models

pytestmark = pytest.mark.gpu

GRAD_TOL = 1e-3        # every parameter gradient and the whole gradient: within 1e-3 relative of fp64 autograd


@pytest.fixture(scope="module")
def gm():
    import gnn_manip_b200
    assert torch.cuda.is_available()
    return gnn_manip_b200


def graph_attr(gm, noise=None, max_neighbours=scene.MAX_NBRS):
    return gm.GraphBoundedMultimaterialControl(scene.CONN_R, scene.make_stats(), scene.CART_IDX, [scene.MAT_IDX],
                                               scene.CTRL_IDX, scene.make_bounds(), noise=noise,
                                               max_neighbours=max_neighbours)


def samples(n, count, n_rigid, h=0.005):
    out = []
    for i in range(count):
        obs, nxt = scene.make_state(n, h=h, seed=20 + i, n_rigid=n_rigid)
        g = torch.Generator().manual_seed(100 + i)
        obs[-1, :n_rigid, 5:8] = torch.randn((n_rigid, 3), generator=g) * 3e-4
        nxt = nxt + torch.randn((n, 3), generator=g) * 5e-5          # a non-trivial target acceleration
        out.append((obs, nxt))
    return out


def oracle_loss_and_grads(ref, feats, batch, sand_only=False, dtype=torch.float32, dpred=None):
    """Loss and parameter gradients by torch autograd.  dpred given: the backward is driven by that upstream gradient
    (a fixed sign pattern of the L1 loss) instead of the oracle's own."""
    nodes, edge_attr, ei, tgt = feats.process_collate(batch)
    ref.zero_grad()
    pred = ref(nodes.to(dtype), edge_attr.to(dtype), ei)
    tgt = tgt.to(dtype)
    if sand_only:
        sel = nodes[:, -4] < 0.5                                     # train_dyn.py:57-58 with use_control
        loss = torch.nn.L1Loss(reduction="sum")(pred[sel], tgt[sel]) / int(sel.sum())
    else:
        loss = torch.nn.L1Loss(reduction="sum")(pred, tgt) / pred.shape[0]
    if dpred is None:
        loss.backward()
    else:
        pred.backward(dpred.to(dtype))
    return loss.detach(), pred.detach(), [p.grad.detach().clone() for p in ref.parameters()]


def double_copy(oracle_lib, ref):
    m = oracle_lib.EncProcDecGNN(25, 4, 3, 128, 2, 10).double()
    m.load_state_dict({k: v.double() for k, v in ref.state_dict().items()})
    return m


def test_flat_parameter_order_matches_module_parameters(gm, oracle_lib):
    ref, mine = seeded_models(gm, oracle_lib, seed=0)
    flat = torch.cat([p.detach().reshape(-1) for p in ref.parameters()])
    assert flat.numel() == 1591299 and torch.equal(flat, mine.flat_parameters())


class FixedMaskReLU(torch.nn.Module):
    """ReLU whose active set is given: x * mask.  With the active sets of OUR forward installed in the oracle model, both
    backward passes differentiate the same linear piece of the network."""

    def __init__(self, mask):
        super().__init__()
        self.mask = mask

    def forward(self, x):
        return x * self.mask


def install_masks(model64, tr, n_nodes, n_edges, eref):
    """Replace every nn.ReLU of the fp64 oracle copy by FixedMaskReLU with the active set saved by the training plan
    (edge rows reordered from receiver-major to the reference's edge order)."""
    def node_mask(which, step):
        return (tr.activation(which, step, n_nodes).cpu() > 0).double()

    def edge_mask(which, step):
        h = tr.activation(which, step, n_edges).cpu()
        out = torch.empty_like(h)
        out[eref] = h
        return (out > 0).double()

    def patch(seq, m1, m2):
        seq[1], seq[3] = FixedMaskReLU(m1), FixedMaskReLU(m2)
    patch(model64.encoder.phi_edge, edge_mask(0, 0), edge_mask(1, 0))
    patch(model64.encoder.phi_node, node_mask(2, 0), node_mask(3, 0))
    for s, blk in enumerate(model64.processor):
        patch(blk.phi_edge, edge_mask(4, s), edge_mask(5, s))
        patch(blk.phi_node, node_mask(6, s), node_mask(7, s))
    patch(model64.decoder, node_mask(8, 0), node_mask(9, 0))


@pytest.mark.parametrize("n,count,h,sand_only", [(400, 2, 0.005, False), (700, 3, 0.012, False), (400, 2, 0.005, True)])
def test_gradients_match_autograd(gm, oracle_lib, n, count, h, sand_only):
    """Every parameter gradient against fp64 torch autograd over the oracle model.

    The network is piecewise linear (ReLU) and the L1 loss has a sign() gradient, so two correct forwards that differ in
    the last bits differentiate neighbouring linear pieces: with a forward error eps a fraction ~eps of the 61 M hidden
    units sits within rounding of its kink, and the gradients differ by ~sqrt(eps) (measured: fp32 autograd against fp64
    autograd 3e-4 .. 8e-4 on these batches; ours, eps = 2e-5, 2e-3 .. 4e-3).  The backward is therefore checked on the
    SAME piece: the active sets and the L1 sign pattern of our forward are installed in the fp64 oracle.  Bar: 1e-3
    relative for every parameter and the whole gradient."""
    ref, mine = seeded_models(gm, oracle_lib, seed=0)
    batch = samples(n, count, n // 3, h)
    feats = oracle_features(oracle_lib)
    loss_ref, pred_ref, free_ref = oracle_loss_and_grads(ref, feats, batch, sand_only)
    b = gm.make_batch(graph_attr(gm), batch, sand_only=sand_only)
    N, E = n * count, b.graph.n_edges
    assert E > 0
    tr = gm.Trainer(mine, N, N * scene.MAX_NBRS)
    loss = tr.forward_backward(b)
    torch.cuda.synchronize()
    assert abs(float(loss) - float(loss_ref)) < 1e-4 * abs(float(loss_ref))
    mine_pred = tr.prediction(N).cpu()
    assert rel_l2(mine_pred, pred_ref) < 1e-4
    nodes, _, _, tgt_ref = feats.process_collate(batch)
    sel = (nodes[:, -4] < 0.5) if sand_only else torch.ones(N, dtype=torch.bool)
    dpred = torch.sign(mine_pred - tgt_ref) * sel[:, None] / int(sel.sum())
    flips = int((torch.sign(mine_pred - tgt_ref) != torch.sign(pred_ref - tgt_ref)).sum())
    model64 = double_copy(oracle_lib, ref)
    install_masks(model64, tr, N, E, b.graph.eref[:E].cpu().long())
    _, pred64, grads_64 = oracle_loss_and_grads(model64, feats, batch, sand_only, torch.float64, dpred)
    assert rel_l2(pred64, pred_ref) < 1e-4                 # the fixed active sets change nothing visible in the forward
    flat = tr.grads.cpu()
    off, report = 0, []
    for (name, _), g in zip(ref.named_parameters(), grads_64):
        mine_g = flat[off: off + g.numel()].reshape(g.shape).double()
        off += g.numel()
        denom = float(g.norm())
        report.append((float((mine_g - g).norm()) / denom if denom > 0 else float(mine_g.norm()), name, denom))
    assert off == flat.numel()
    report.sort(reverse=True)
    flat64 = torch.cat([g.reshape(-1) for g in grads_64])
    total = rel_l2(flat, flat64)
    free = rel_l2(torch.cat([g.reshape(-1) for g in free_ref]), flat64)
    print(f"n={n}x{count} E={E} loss {float(loss):.6f} vs {float(loss_ref):.6f}; whole gradient rel_l2 {total:.2e}; "
          f"largest per-parameter errors {[(f'{e:.1e}', k) for e, k, _ in report[:4]]}; L1 signs differing between the "
          f"forwards {flips} of {mine_pred.numel()}; free-running fp32 autograd against the same yardstick {free:.2e}")
    for err, name, _ in report:
        assert err < GRAD_TOL, f"gradient of {name}: relative error {err:.3e}"
    assert total < GRAD_TOL
    # bit-reproducible: the weight-gradient reductions run in a fixed order
    first = tr.grads.clone()
    tr.forward_backward(b)
    assert torch.equal(first, tr.grads)
    tr.close()


def test_training_steps_match_torch_adam(gm, oracle_lib):
    ref, mine = seeded_models(gm, oracle_lib, seed=1)
    feats = oracle_features(oracle_lib)
    batches = [samples(300, 2, 100)[::-1], samples(300, 2, 100)]
    opt = torch.optim.Adam(ref.parameters(), lr=1e-4)                 # train_dyn.py:98
    tr = gm.Trainer(mine, 600, 600 * scene.MAX_NBRS, lr=1e-4)
    ga = graph_attr(gm)
    for step in range(3):
        batch = batches[step % 2]
        loss_ref, _, _ = oracle_loss_and_grads(ref, feats, batch)
        opt.step()
        loss = tr.loss_batch(gm.make_batch(ga, batch))
        assert abs(float(loss) - float(loss_ref)) < 2e-4 * abs(float(loss_ref)), f"step {step}"
    want = torch.cat([p.detach().reshape(-1) for p in ref.parameters()])
    got = tr.params.cpu()
    # Adam divides by sqrt(v): an ill-conditioned, near-zero gradient entry may move its weight the other way, so the
    # comparison is on the distribution of the differences (three steps of lr 1e-4 move a weight by at most 3e-4)
    diff = (got - want).abs()
    q = torch.quantile(diff[torch.randperm(diff.numel())[:200000]], torch.tensor([0.5, 0.99, 0.999]))
    print("Adam 3 steps: |param diff| median / 99% / 99.9% / max:", [f"{float(v):.2e}" for v in q], f"{float(diff.max()):.2e}")
    assert float(q[0]) < 2e-6 and float(q[1]) < 5e-5 and float(q[2]) < 3e-4 and float(diff.max()) < 6.5e-4
    moved = float((got - mine.flat_parameters()).abs().max())
    assert moved > 1e-4
    # test=True leaves the parameters alone; export writes the state dict layout back
    before = tr.params.clone()
    tr.loss_batch(gm.make_batch(ga, batches[0]), test=True)
    assert torch.equal(before, tr.params)
    tr.export_to(mine)
    assert torch.equal(mine.flat_parameters(), tr.params.cpu()) and len(mine.state_dict()) == 342
    with torch.no_grad():                                             # the inference kernels see the trained weights
        trained = oracle_lib.EncProcDecGNN(25, 4, 3, 128, 2, 10)
        trained.load_state_dict(mine.state_dict(), strict=True)
        nodes, edge_attr, ei, _ = feats.process_collate(batches[0])
        assert rel_l2(mine(nodes, edge_attr, ei), trained(nodes, edge_attr, ei)) < 1e-3
    tr.close()


def test_batch_assembly_matches_oracle(gm, oracle_lib):
    batch = samples(500, 3, 150)
    nodes, edge_attr, ei, tgt = oracle_features(oracle_lib).process_collate(batch)
    b = gm.make_batch(graph_attr(gm), batch)
    assert torch.equal(b.nodes.cpu(), nodes) and torch.equal(b.target.cpu(), tgt)
    assert torch.equal(b.graph.edge_index().cpu(), ei)
    sand = gm.make_batch(graph_attr(gm), batch, sand_only=True)
    assert torch.equal(sand.row_mask.cpu().bool(), nodes[:, -4] < 0.5)


def test_random_walk_noise(gm):
    k, n, std = 6, 20000, 3e-4
    pos = torch.zeros((k, n, 3))
    noise, z = gm.random_walk_noise(pos, std, seed=7, return_samples=True)
    assert noise.shape == (k, n, 3) and z.shape == (k - 1, n, 3)
    assert torch.equal(noise[0], torch.zeros((n, 3)))
    # structure (utils.py:104-112): two cumulative sums of the per-step velocity noise, fp32, in time order
    want = torch.zeros_like(noise)
    vel = torch.zeros((n, 3))
    acc = torch.zeros((n, 3))
    for t in range(k - 1):
        vel = vel + z[t]
        acc = acc + vel
        want[t + 1] = acc
    assert torch.equal(noise, want)
    # distribution (utils.py:100-103): per-step std = noise_std / sqrt(k - 1); the velocity noise of the last step has
    # std = noise_std; draws are independent across particles, axes and steps
    assert abs(float(z.std()) / (std / (k - 1) ** 0.5) - 1) < 0.01 and abs(float(z.mean())) < 3 * std / (z.numel() ** 0.5)
    assert abs(float(z.sum(0).std()) / std - 1) < 0.02
    c = np.corrcoef(z[0, :, 0].numpy(), z[1, :, 0].numpy())[0, 1]
    assert abs(c) < 0.03
    again = gm.random_walk_noise(pos, std, seed=7)
    other = gm.random_walk_noise(pos, std, seed=8)
    assert torch.equal(again, noise) and not torch.equal(other, noise)


def test_process_noisy_path(gm, oracle_lib):
    """collate_utils.py:169-193: noisy observation -> features, graph of the noisy last frame with the CLASS's
    max_neighbours, target from the noisy window and the noise-shifted next position."""
    obs, nxt = samples(400, 1, 120)[0]
    ga = graph_attr(gm, noise=3e-4, max_neighbours=12)
    torch.manual_seed(0)
    nodes, edge_attr, s, r, tgt = ga.process(obs, nxt)
    assert nodes.shape == (400, 25) and edge_attr.shape[1] == 4 and tgt.shape == (400, 3)
    assert int(torch.bincount(s).max()) <= 12
    clean = graph_attr(gm, noise=None).process(obs, nxt)
    assert not torch.equal(nodes[:, :15], clean[0][:, :15])           # velocities carry the noise
    assert torch.equal(nodes[:, 21], clean[0][:, 21])                 # the material column does not
    # device batch path: same construction, seeded -> reproducible; noise-free graph_attr reproduces the clean batch
    b1 = gm.make_batch(ga, [(obs, nxt)], seed=5)
    b2 = gm.make_batch(ga, [(obs, nxt)], seed=5)
    assert torch.equal(b1.nodes, b2.nodes) and torch.equal(b1.target, b2.target)
    # consistency of the noisy target: acceleration of (noisy window, noisy next position)
    feats = oracle_features(oracle_lib)
    L = gm._lib.lib()
    import ctypes
    o = obs.cuda().contiguous()
    t = nxt.cuda().contiguous()
    no, nt = torch.empty_like(o), torch.empty_like(t)
    phys = ga.phys()
    gm._lib.check(L.gm_random_walk_noise(gm._lib.ptr(o), 6, 400, 8, ctypes.byref(phys), gm._lib.ptr(t), 3e-4,
                                         ctypes.c_uint64(5), gm._lib.ptr(no), gm._lib.ptr(nt), None,
                                         gm._lib.stream_ptr()), "noise")
    assert torch.equal(no[..., :2], o[..., :2]) and torch.equal(no[..., 5:], o[..., 5:])
    assert torch.allclose(nt - t, no[-1, :, 2:5] - o[-1, :, 2:5], rtol=0, atol=2e-7)   # tgt += noise[-1] (collate_utils.py:181)
    assert torch.equal(b1.target.cpu(), feats.compute_target(no.cpu(), nt.cpu()))
    assert torch.equal(b1.nodes.cpu(), feats.compute_nodes(no.cpu()))
