"""GPU suite (-m gpu): the kernels the training step is built from (rows E(2) / N3), each against a plain PyTorch
fp32/fp64 reference of the same op on the same device, called through the C ABI."""
import ctypes

import numpy as np
import pytest
import torch

from conftest import rel_l2

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gm():
    import gnn_manip_b200
    assert torch.cuda.is_available()
    return gnn_manip_b200


def _chunk():
    buf = torch.zeros(65536 + 1024, dtype=torch.uint8, device="cuda")
    off = (-buf.data_ptr()) % 1024
    return buf, buf.data_ptr() + off


def pack(gm, w, ld, col0, transpose):
    buf, addr = _chunk()
    job = gm._lib.gm_pack_job()
    job.w, job.ld, job.col0, job.transpose, job.chunk = w.data_ptr(), ld, col0, transpose, addr
    jobs = torch.frombuffer(bytearray(bytes(job)), dtype=torch.uint8).cuda()
    gm._lib.check(gm._lib.lib().gm_train_pack(gm._lib.ptr(jobs), 1, gm._lib.stream_ptr()), "gm_train_pack")
    torch.cuda.synchronize()
    return buf, addr


@pytest.mark.parametrize("M", [1, 128, 300, 148 * 128 * 2 + 77])
def test_gemm_forward_and_epilogues(gm, M):
    L, lib = gm._lib.lib(), gm._lib
    g = torch.Generator(device="cuda").manual_seed(M)
    x = torch.randn((M + 5, 128), device="cuda", generator=g)
    w = torch.randn((128, 384), device="cuda", generator=g) / 11.0
    bias = torch.randn((128,), device="cuda", generator=g)
    add1 = torch.randn((50, 128), device="cuda", generator=g)
    add2 = torch.randn((M, 128), device="cuda", generator=g)
    idx_in = torch.randint(0, M + 5, (M,), device="cuda", generator=g, dtype=torch.int32)
    idx1 = torch.randint(0, 50, (M,), device="cuda", generator=g, dtype=torch.int32)
    keep, addr = pack(gm, w, 384, 128, 0)
    out = torch.full((M, 128), float("nan"), device="cuda")
    lib.check(L.gm_train_gemm(lib.ptr(x), lib.ptr(idx_in), ctypes.c_void_p(addr), lib.ptr(bias), lib.ptr(add1),
                              lib.ptr(idx1), lib.ptr(add2), None, None, 1, 0, M, None, lib.ptr(out), lib.stream_ptr()),
              "gm_train_gemm")
    ref = torch.relu(x[idx_in.long()].double() @ w[:, 128:256].double().T + bias.double() + add1[idx1.long()].double()
                     + add2.double())
    assert rel_l2(out.cpu(), ref.cpu()) < 2e-5
    # dgrad flavour: transposed pack, ReLU mask, accumulate into an existing gradient, device-side row count
    keep2, addr2 = pack(gm, w, 384, 256, 1)
    mask = torch.randn((M, 128), device="cuda", generator=g)
    prev = torch.randn((M, 128), device="cuda", generator=g)
    out2 = prev.clone()
    m_dev = torch.tensor([max(M - 3, 1)], dtype=torch.int32, device="cuda")
    lib.check(L.gm_train_gemm(lib.ptr(x), None, ctypes.c_void_p(addr2), None, None, None, None, None, lib.ptr(mask), 0,
                              1, M, lib.ptr(m_dev), lib.ptr(out2), lib.stream_ptr()), "gm_train_gemm")
    live = int(m_dev)
    ref2 = prev.double().clone()
    ref2[:live] += (x[:M].double() @ w[:, 256:384].double())[:live] * (mask[:live] > 0)
    assert rel_l2(out2.cpu(), ref2.cpu()) < 2e-5
    assert torch.equal(out2[live:], prev[live:])


@pytest.mark.parametrize("M", [1, 64, 1000, 148 * 64 * 3 + 17])
def test_wgrad_matches_torch(gm, M):
    L, lib = gm._lib.lib(), gm._lib
    g = torch.Generator(device="cuda").manual_seed(M + 1)
    A = torch.randn((M, 128), device="cuda", generator=g)
    B = torch.randn((M, 128), device="cuda", generator=g)
    dW = torch.randn((128, 384), device="cuda", generator=g)
    db = torch.randn((128,), device="cuda", generator=g)
    dW0, db0 = dW.clone(), db.clone()
    part = torch.empty((L.gm_train_wgrad_partial_bytes(),), dtype=torch.uint8, device="cuda")
    lib.check(L.gm_train_wgrad(lib.ptr(A), lib.ptr(B), M, None, lib.ptr(dW), 384, 128, lib.ptr(db), lib.ptr(part),
                               lib.stream_ptr()), "gm_train_wgrad")
    want = dW0.double().clone()
    want[:, 128:256] += A.double().T @ B.double()
    assert rel_l2((dW - dW0).cpu()[:, 128:256], (want - dW0.double()).cpu()[:, 128:256]) < 2e-5
    assert torch.equal(dW[:, :128], dW0[:, :128]) and torch.equal(dW[:, 256:], dW0[:, 256:])
    assert rel_l2((db - db0).cpu(), A.double().sum(0).cpu()) < 1e-5
    # deterministic: a second run from the same start gives the same bits
    dW1, db1 = dW0.clone(), db0.clone()
    lib.check(L.gm_train_wgrad(lib.ptr(A), lib.ptr(B), M, None, lib.ptr(dW1), 384, 128, lib.ptr(db1), lib.ptr(part),
                               lib.stream_ptr()), "gm_train_wgrad")
    assert torch.equal(dW1, dW) and torch.equal(db1, db)


@pytest.mark.parametrize("M", [1, 77, 5000])
def test_layernorm_forward_backward(gm, M):
    L, lib = gm._lib.lib(), gm._lib
    g = torch.Generator(device="cuda").manual_seed(M + 2)
    z = (torch.randn((M, 128), device="cuda", generator=g) * 3 + 1).requires_grad_(True)
    gamma = torch.randn((128,), device="cuda", generator=g).requires_grad_(True)
    beta = torch.randn((128,), device="cuda", generator=g).requires_grad_(True)
    res = torch.randn((M, 128), device="cuda", generator=g)
    xhat, y, out = (torch.empty((M, 128), device="cuda") for _ in range(3))
    rstd = torch.empty((M,), device="cuda")
    lib.check(L.gm_train_layernorm_fwd(lib.ptr(z), lib.ptr(gamma), lib.ptr(beta), M, lib.ptr(xhat), lib.ptr(rstd),
                                       lib.ptr(y), lib.ptr(res), lib.ptr(out), lib.stream_ptr()), "ln_fwd")
    yr = torch.nn.functional.layer_norm(z, (128,), gamma, beta, 1e-5)
    assert rel_l2(y.cpu(), yr.detach().cpu()) < 1e-6 and rel_l2(out.cpu(), (yr + res).detach().cpu()) < 1e-6
    d1 = torch.randn((M, 128), device="cuda", generator=g)
    d2 = torch.randn((9, 128), device="cuda", generator=g)
    idx2 = torch.randint(0, 9, (M,), device="cuda", generator=g, dtype=torch.int32)
    dy = d1 + d2[idx2.long()]
    yr.backward(dy)
    dz = torch.empty((M, 128), device="cuda")
    dg, dbt = torch.zeros((128,), device="cuda"), torch.zeros((128,), device="cuda")
    part = torch.empty((L.gm_train_layernorm_partial_bytes(M),), dtype=torch.uint8, device="cuda")
    lib.check(L.gm_train_layernorm_bwd(lib.ptr(d1), lib.ptr(d2), lib.ptr(idx2), lib.ptr(xhat), lib.ptr(rstd),
                                       lib.ptr(gamma), M, lib.ptr(dz), lib.ptr(dg), lib.ptr(dbt), lib.ptr(part),
                                       lib.stream_ptr()), "ln_bwd")
    assert rel_l2(dz.cpu(), z.grad.cpu()) < 2e-5
    assert rel_l2(dg.cpu(), gamma.grad.cpu()) < 1e-5 and rel_l2(dbt.cpu(), beta.grad.cpu()) < 1e-5


@pytest.mark.parametrize("K,O,relu", [(4, 128, 1), (25, 128, 1), (128, 3, 0)])
def test_small_layers_forward_backward(gm, K, O, relu):
    L, lib = gm._lib.lib(), gm._lib
    M = 1000
    g = torch.Generator(device="cuda").manual_seed(K * O)
    x = torch.randn((M, K), device="cuda", generator=g).requires_grad_(True)
    w = (torch.randn((O, K), device="cuda", generator=g) / K ** 0.5).requires_grad_(True)
    b = torch.randn((O,), device="cuda", generator=g).requires_grad_(True)
    h = torch.empty((M, O), device="cuda")
    lib.check(L.gm_train_small_fwd(lib.ptr(x), K, lib.ptr(w), lib.ptr(b), O, M, relu, lib.ptr(h), lib.stream_ptr()),
              "small_fwd")
    hr = torch.nn.functional.linear(x, w, b)
    hr = torch.relu(hr) if relu else hr
    assert rel_l2(h.cpu(), hr.detach().cpu()) < 1e-6
    dh = torch.randn((M, O), device="cuda", generator=g)
    hr.backward(dh)
    dW, db, dx = torch.zeros((O, K), device="cuda"), torch.zeros((O,), device="cuda"), torch.empty((M, K), device="cuda")
    part = torch.empty((L.gm_train_small_partial_bytes(M, K, O),), dtype=torch.uint8, device="cuda")
    lib.check(L.gm_train_small_bwd(lib.ptr(x), K, lib.ptr(w), O, lib.ptr(dh), lib.ptr(h) if relu else None, M,
                                   lib.ptr(dW), lib.ptr(db), lib.ptr(dx), lib.ptr(part), lib.stream_ptr()), "small_bwd")
    assert rel_l2(dW.cpu(), w.grad.cpu()) < 1e-5 and rel_l2(db.cpu(), b.grad.cpu()) < 1e-5
    assert rel_l2(dx.cpu(), x.grad.cpu()) < 1e-5


def test_l1_loss_and_adam(gm):
    L, lib = gm._lib.lib(), gm._lib
    M = 3000
    g = torch.Generator(device="cuda").manual_seed(5)
    pred = torch.randn((M, 3), device="cuda", generator=g).requires_grad_(True)
    tgt = torch.randn((M, 3), device="cuda", generator=g)
    tgt[5] = pred[5].detach()                                  # |0| has gradient 0 in torch
    rows = torch.rand((M,), device="cuda", generator=g) < 0.6
    scratch = torch.empty((2 + 2 * ((M + 255) // 256),), device="cuda")
    for mask in (None, rows):
        loss, dpred = torch.zeros((1,), device="cuda"), torch.empty((M, 3), device="cuda")
        m8 = None if mask is None else mask.to(torch.uint8)
        lib.check(L.gm_train_l1_loss(lib.ptr(pred), lib.ptr(tgt), lib.ptr(m8), M, 3, lib.ptr(loss), lib.ptr(dpred),
                                     lib.ptr(scratch), lib.stream_ptr()), "l1")
        sel = slice(None) if mask is None else mask
        ref = torch.nn.L1Loss(reduction="sum")(pred[sel], tgt[sel]) / pred[sel].shape[0]    # train_dyn.py:56-64
        pred.grad = None
        ref.backward()
        assert abs(float(loss) - float(ref)) < 1e-5 * abs(float(ref))
        assert torch.allclose(dpred, pred.grad, rtol=1e-6, atol=0)
    # Adam: 3 steps against torch.optim.Adam
    p0 = torch.randn((10001,), device="cuda", generator=g)
    p_ref = p0.clone().requires_grad_(True)
    opt = torch.optim.Adam([p_ref], lr=1e-4)
    p, m, v = p0.clone(), torch.zeros_like(p0), torch.zeros_like(p0)
    for t in range(1, 4):
        grad = torch.randn((10001,), device="cuda", generator=g)
        p_ref.grad = grad.clone()
        opt.step()
        lib.check(L.gm_adam_step(lib.ptr(p), lib.ptr(grad), lib.ptr(m), lib.ptr(v), p.numel(), 1e-4, 0.9, 0.999, 1e-8,
                                 t, lib.stream_ptr()), "adam")
        assert float((p - p_ref.detach()).abs().max()) < 2e-7
