"""GPU suite (-m gpu), collected last on purpose: gm_train_layernorm_bwd_static -- the LayerNorm backward whose scratch
is a file-scope device array of the library instead of a caller-provided buffer (no counterpart in the reference; added for
callers that cannot allocate).  Its results must equal the existing entry point's, on two calls in a row (the shared
array keeps the previous call's contents), and sizes beyond the stated maximum must be rejected."""
import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_ROWS = 65536      # GM_LN_STATIC_MAX_ROWS in include/gnn_manip_b200.h


@pytest.fixture(scope="module")
def gm():
    import gnn_manip_b200
    assert torch.cuda.is_available()
    return gnn_manip_b200


def _inputs(M, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    d1 = torch.randn((M, 128), device="cuda", generator=g)
    d2 = torch.randn((9, 128), device="cuda", generator=g)
    idx2 = torch.randint(0, 9, (M,), device="cuda", generator=g, dtype=torch.int32)
    xhat = torch.randn((M, 128), device="cuda", generator=g)
    rstd = torch.rand((M,), device="cuda", generator=g) + 0.5
    gamma = torch.randn((128,), device="cuda", generator=g)
    return d1, d2, idx2, xhat, rstd, gamma


def _existing(gm, inp, M):
    L, lib = gm._lib.lib(), gm._lib
    d1, d2, idx2, xhat, rstd, gamma = inp
    dz = torch.empty((M, 128), device="cuda")
    dg, db = torch.zeros((128,), device="cuda"), torch.zeros((128,), device="cuda")
    part = torch.empty((L.gm_train_layernorm_partial_bytes(M),), dtype=torch.uint8, device="cuda")
    lib.check(L.gm_train_layernorm_bwd(lib.ptr(d1), lib.ptr(d2), lib.ptr(idx2), lib.ptr(xhat), lib.ptr(rstd),
                                       lib.ptr(gamma), M, lib.ptr(dz), lib.ptr(dg), lib.ptr(db), lib.ptr(part),
                                       lib.stream_ptr()), "ln_bwd")
    torch.cuda.synchronize()
    return dz, dg, db


def _static(gm, inp, M, dz, dg, db):
    L, lib = gm._lib.lib(), gm._lib
    d1, d2, idx2, xhat, rstd, gamma = inp
    lib.check(L.gm_train_layernorm_bwd_static(lib.ptr(d1), lib.ptr(d2), lib.ptr(idx2), lib.ptr(xhat), lib.ptr(rstd),
                                              lib.ptr(gamma), M, lib.ptr(dz), lib.ptr(dg), lib.ptr(db),
                                              lib.stream_ptr()), "ln_bwd_static")
    torch.cuda.synchronize()


def test_static_scratch_layernorm_backward_matches_existing_entry_point(gm):
    # two calls in a row of different sizes: the second finds the first one's partials in the shared array (a larger
    # call first, so that stale slices beyond the second call's block count exist)
    for M, seed in ((5000, 11), (777, 12), (MAX_ROWS, 13)):
        inp = _inputs(M, seed)
        dz_ref, dg_ref, db_ref = _existing(gm, inp, M)
        # outputs pre-allocated: nothing is allocated between here and the result (the entry point itself never allocates)
        dz = torch.empty((M, 128), device="cuda")
        dg, db = torch.zeros((128,), device="cuda"), torch.zeros((128,), device="cuda")
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated()
        _static(gm, inp, M, dz, dg, db)
        assert torch.cuda.memory_allocated() == before
        # same kernels, same fixed-order reduction: bit-identical (the tolerance of the existing test, 1e-5, holds a fortiori)
        assert torch.equal(dz, dz_ref) and torch.equal(dg, dg_ref) and torch.equal(db, db_ref)
        # second call on the same inputs accumulates once more into dgamma / dbeta, exactly like the existing entry point
        _static(gm, inp, M, dz, dg, db)
        assert torch.equal(dz, dz_ref)
        assert torch.allclose(dg, 2 * dg_ref, rtol=1e-6, atol=0) and torch.allclose(db, 2 * db_ref, rtol=1e-6, atol=0)


def test_static_scratch_rejects_sizes_beyond_the_maximum(gm):
    L, lib = gm._lib.lib(), gm._lib
    M = 128
    inp = _inputs(M, 5)
    d1, d2, idx2, xhat, rstd, gamma = inp
    dz = torch.empty((M, 128), device="cuda")
    dg, db = torch.zeros((128,), device="cuda"), torch.zeros((128,), device="cuda")
    rc = L.gm_train_layernorm_bwd_static(lib.ptr(d1), lib.ptr(d2), lib.ptr(idx2), lib.ptr(xhat), lib.ptr(rstd),
                                         lib.ptr(gamma), MAX_ROWS + 1, lib.ptr(dz), lib.ptr(dg), lib.ptr(db),
                                         lib.stream_ptr())
    assert rc == 1                                         # GM_STATUS_INVALID
    msg = L.gm_last_error().decode()
    assert str(MAX_ROWS) in msg and "static scratch" in msg
    with pytest.raises(gm._lib.GmError):
        lib.check(rc, "ln_bwd_static")
    torch.cuda.synchronize()
    assert float(dg.abs().sum()) == 0.0                    # nothing was launched
