"""CPU suite for rows N1 / N2: the oracle's trajectory parametrisations and loss terms against the goldens recorded
from the unmodified reference classes (tests/golden/traj_param.npz), the numpy Sinkhorn statement's invariants, and the
host CMA-ES ask/tell loop (pure host logic of the product: no kernel involved)."""
import numpy as np
import pytest

from conftest import load_golden
from oracle import scene


def _inc(restate):
    return restate.TrajectoryParam(scene.RX_INIT_DEG, scene.TY_INIT, 0.05, 3.0, 0.002)


def _pchip(restate, total_steps, traj_points):
    return restate.PchipParam(scene.RX_INIT_DEG, scene.TY_INIT, 0.05, 3.0, 0.002, total_steps, traj_points)


def test_increment_parametrisation_matches_reference_golden(oracle_lib):
    g = load_golden("traj_param")
    st = load_golden_sample()
    p = _inc(oracle_lib)
    assert np.array_equal(p.x0_from_sample(st), g["inc_x"][0])
    assert p.scale_ty == float(g["inc_scale_ty"])
    for i, x in enumerate(g["inc_x"]):
        rot, ty = p.interpolate(x)
        assert np.array_equal(np.array(rot), g["inc_rot"][i]) and np.array_equal(np.array(ty), g["inc_ty"][i])
        assert np.allclose(oracle_lib.increments_terms(p, x), g["inc_terms"][i], rtol=1e-12, atol=0)


def load_golden_sample():
    import os
    from conftest import GOLDEN
    return np.load(os.path.join(GOLDEN, "sample_traj.npy"))


@pytest.mark.parametrize("key,total,points", [("p150", 150, 50), ("p300", 300, 10)])
def test_pchip_parametrisation_matches_reference_golden(oracle_lib, key, total, points):
    g = load_golden("traj_param")
    st = load_golden_sample()
    p = _pchip(oracle_lib, total, points)
    assert np.array_equal(p.x0_from_sample(st[: total + 1]), g[f"{key}_x"][0])
    for i, x in enumerate(g[f"{key}_x"]):
        rot, ty = p.interpolate(x)
        assert np.array_equal(rot, g[f"{key}_rot"][i]), f"{key} candidate {i}: rotation curve differs from scipy pchip"
        assert np.array_equal(ty, g[f"{key}_ty"][i])
        assert np.allclose(p.terms(x), g[f"{key}_terms"][i], rtol=1e-12, atol=0)
        if key == "p150":
            assert np.array_equal(p.ineq_constraint(x), g["p150_ineq"][i])


def test_rigid_trajectory_from_parameters_matches_reference_golden(oracle_lib):
    import torch
    g = load_golden("traj_param")
    p = _pchip(oracle_lib, 150, 50)
    rot, ty = p.interpolate(g["p150_x"][3])
    obs = torch.from_numpy(g["p150_obs0"])
    rigid = obs[-1, obs[-1, :, scene.MAT_IDX] == 1][:, scene.CART_IDX]
    traj = oracle_lib.rigid_body_trajectory(rot, ty, 150, scene.TY_INIT, rigid)
    assert float((traj - torch.from_numpy(g["p150_rigid_x3"])).abs().max()) < 1e-7


def test_sinkhorn_statement_invariants(oracle_lib):
    """geomloss is absent (parity unpinned): the numpy statement is held to the properties of the debiased divergence --
    S(x, x) = 0, translation by t costs |t|^2 / 2 exactly, symmetry, positivity."""
    rng = np.random.RandomState(0)
    x = 0.5 + 0.04 * rng.randn(300, 3)
    y = 0.5 + 0.04 * rng.randn(260, 3) + np.array([0.03, 0.0, -0.01])
    s = oracle_lib.sinkhorn_divergence
    assert abs(s(x, x)) < 1e-15
    assert abs(s(x, x + np.array([0.1, 0.0, 0.0])) - 0.005) < 1e-8
    a, b = s(x, y), s(y, x)
    assert a > 0 and abs(a - b) < 1e-12 * max(1.0, abs(a))
    sched = oracle_lib.sinkhorn_epsilon_schedule(0.4, 0.05, 0.5)
    assert sched[0] == 0.4 ** 2 and sched[-1] == 0.05 ** 2 and all(sched[i] >= sched[i + 1] for i in range(len(sched) - 1))
    assert len(sched) == 2 + int(np.ceil(np.log(0.4 / 0.05) / np.log(2.0)))


def test_cma_es_minimises_and_is_seeded():
    from gnn_manip_b200 import cma_es

    def sphere(X):
        return ((X - 0.7) ** 2).sum(axis=1)

    def rosen(X):
        return (100.0 * (X[:, 1:] - X[:, :-1] ** 2) ** 2 + (1 - X[:, :-1]) ** 2).sum(axis=1)

    xb, es = cma_es.fmin_batch(sphere, np.zeros(6), 0.5, {"popsize": 16, "seed": 1234, "maxiter": 150})
    assert es.best_f < 1e-12 and np.allclose(xb, 0.7, atol=1e-5)
    xb2, es2 = cma_es.fmin_batch(sphere, np.zeros(6), 0.5, {"popsize": 16, "seed": 1234, "maxiter": 150})
    assert np.array_equal(xb, xb2) and es.history == es2.history
    xb, es = cma_es.fmin_batch(rosen, np.zeros(4), 0.5, {"popsize": 16, "seed": 3, "maxiter": 600})
    assert es.best_f < 1e-8
    # box bounds: the optimum 0.7 lies outside [-0.5, 0.5] -> the best feasible point is the corner 0.5
    xb, es = cma_es.fmin_batch(sphere, np.zeros(3), 0.3, {"popsize": 12, "seed": 1, "maxiter": 200, "bounds": [-0.5, 0.5]})
    assert np.all(np.abs(xb) <= 0.5) and np.allclose(xb, 0.5, atol=1e-3)
    # inequality constraint x_0 <= 0.2 through the augmented Lagrangian
    xb, es = cma_es.fmin_batch(sphere, np.zeros(3), 0.3, {"popsize": 12, "seed": 2, "maxiter": 300},
                               constraints_batch=lambda X: X[:, :1] - 0.2)
    assert abs(xb[0] - 0.2) < 2e-2 and np.allclose(xb[1:], 0.7, atol=2e-2)
    es = cma_es.CMAES(np.zeros(5), 0.5, popsize=64)
    assert es.ask().shape == (64, 5)
    with pytest.raises(ValueError):
        es.tell(np.zeros(3))
