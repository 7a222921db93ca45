"""CPU suite, build container only: the oracle restatement (oracle/restate.py) against the UNMODIFIED reference
imported live from /root/reference through oracle/ref_shims.py.  Skipped where the reference checkout is absent
(the GPU box); there the committed goldens (tests/golden, written from the same reference by oracle/make_golden.py)
pin the oracle instead (tests/test_oracle.py)."""
import contextlib
import io

import numpy as np
import pytest
import torch

from conftest import rel_l2
from oracle import ref_shims, scene

pytestmark = pytest.mark.skipif(not ref_shims.reference_available(), reason="/root/reference is not present")


@pytest.fixture(scope="module")
def ref():
    return ref_shims.load()


def _ref_graph_attr(ref):
    return ref.collate.GraphBoundedMultimaterialControl(
        scene.CONN_R, scene.make_stats(), scene.CART_IDX, [scene.MAT_IDX], scene.CTRL_IDX, scene.make_bounds(),
        noise=None, max_neighbours=scene.MAX_NBRS)


def _features(restate):
    return restate.GraphFeatures(scene.CONN_R, scene.make_stats(), scene.make_bounds(), scene.CART_IDX,
                                 scene.MAT_IDX, scene.CTRL_IDX, scene.MAX_NBRS)


@pytest.mark.parametrize("n,h,K", [(900, 0.005, 20), (700, 0.012, 20), (500, 0.004, 7)])
def test_graph_live(ref, oracle_lib, n, h, K):
    pos, _ = scene.lattice_positions(n, h, seed=n)
    s0, r0 = ref.utils.get_connectivity(pos, scene.CONN_R, K)
    for fn in (oracle_lib.get_connectivity, oracle_lib.get_connectivity_bruteforce):
        s1, r1 = fn(pos, scene.CONN_R, K)
        assert torch.equal(s0, s1) and torch.equal(r0, r1)


def test_features_model_integrator_live(ref, oracle_lib):
    n = 500
    obs, nxt = scene.make_state(n, seed=21, n_rigid=150)
    obs[-1, :150, 5:8] = torch.randn((150, 3), generator=torch.Generator().manual_seed(2)) * 3e-4
    n0, e0, ei0, t0 = _ref_graph_attr(ref).process_collate([(obs, nxt)])
    n1, e1, ei1, t1 = _features(oracle_lib).process_collate([(obs, nxt)])
    assert torch.equal(ei0, ei1) and torch.equal(n0, n1) and torch.equal(e0, e1) and torch.equal(t0, t1)
    torch.manual_seed(3)
    with contextlib.redirect_stdout(io.StringIO()):
        m0 = ref.EncProcDecGNN(25, 4, 3, 128, 2, 10)
    torch.manual_seed(3)
    m1 = oracle_lib.EncProcDecGNN(25, 4, 3, 128, 2, 10)
    sd0, sd1 = m0.state_dict(), m1.state_dict()
    assert list(sd0.keys()) == list(sd1.keys()) and len(sd0) == 342
    assert all(torch.equal(sd0[k], sd1[k]) for k in sd0)
    with torch.no_grad():
        p0, p1 = m0.forward(n0, e0, ei0), m1(n1, e1, ei1)
    assert rel_l2(p1, p0) < 2e-6
    stats = scene.make_stats()
    x0 = ref.rollout.get_position_from_prediction(stats, scene.CART_IDX, p0, obs)
    x1 = oracle_lib.get_position_from_prediction(stats, scene.CART_IDX, p0, obs)
    assert torch.equal(x0, x1)


def test_rollout_loop_live(ref, oracle_lib):
    """The reference's cma_objective loop (traj_utils.py:114-159) against oracle.rollout, including steps past the end
    of the control trajectory (the quirk branch :128-129)."""
    n, n_rigid, t_len, horizon = 300, 90, 3, 5
    obs, nxt = scene.make_state(n, seed=9, n_rigid=n_rigid)
    torch.manual_seed(0)
    with contextlib.redirect_stdout(io.StringIO()):
        model = ref.EncProcDecGNN(25, 4, 3, 128, 2, 10)
    rot = np.deg2rad(180.0 + 1.5 * np.arange(t_len))
    ty = -1e-4 * np.arange(t_len)
    final = {}

    class Solver(ref.traj.CMAESolver):
        def get_rigid_body_trajectory_from_diff(self, x, demo=False):      # a trajectory shorter than the horizon
            traj = torch.stack([self.compute_particles_tmatrix(rot[i], ty[i]) for i in range(t_len)])
            return traj, np.zeros((self.horizon, 2))

        def compute_loss(self, end_position, actions, cup_states=None, coffee_states=None, x=None):
            final["end"] = end_position.clone()
            return 0.0, None, None, None, None, None

    with contextlib.redirect_stdout(io.StringIO()):
        solver = Solver(model, _ref_graph_attr(ref), (obs.clone(), nxt.clone()), scene.RX_INIT_DEG, scene.TY_INIT,
                        0.05, 1.0, 1.0, 1.0, 1.0, 1.0, 1.0, "cpu", total_steps=horizon, traj_points=1)
    solver.cma_objective(np.zeros(2))
    torch.manual_seed(0)
    mine = oracle_lib.EncProcDecGNN(25, 4, 3, 128, 2, 10)
    feats = _features(oracle_lib)
    traj = oracle_lib.rigid_body_trajectory(rot, ty, t_len, scene.TY_INIT, obs[-1, :n_rigid][:, scene.CART_IDX])
    out = oracle_lib.rollout(mine, feats, (obs, None), traj, feats.stats, horizon)
    assert float((out[-1, n_rigid:, 2:5] - final["end"]).abs().max()) < 1e-7
