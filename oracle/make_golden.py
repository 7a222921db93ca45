"""TEST INFRASTRUCTURE ONLY -- writes tests/golden/*.npz by running the UNMODIFIED reference.

Run in the build container only (needs /root/reference):  python -m oracle.make_golden
The reference's own files are imported through oracle/ref_shims.py (import shims for absent third-party leaves,
SURVEY.md section 8 row C); nothing of the reference is copied or edited.  The rollout golden runs the reference's
CMAESolver.cma_objective loop itself (gnn_manip/utils/traj_utils.py:114-159) through a subclass that only supplies
the two abstract hooks (interpolate_trajectory -> a given control curve, compute_loss -> capture the end state).

Fixtures (all small, committed):
  graph_dense.npz / graph_sparse.npz   pos -> senders, receivers       (get_connectivity, utils.py:64-93)
  step.npz     obs -> nodes, edge_attr, edge_index, pred, next_pos     (process_collate + EncProcDecGNN.forward +
                                                                        get_position_from_prediction)
  rollout.npz  obs0, rot, ty -> per-step pred, final obs               (cma_objective loop)
  blocks.npz   the 5-node / 6-edge fixture of deps/torch-graphnet/test/gnn_test.py:81-133 through the reference's
               EdgeModel / NodeModel / InteractionNetwork with seed-123 Linear layers
"""
import contextlib
import io
import os

import numpy as np
import torch

from . import ref_shims, scene

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def weights_checksum(sd):
    flat = torch.cat([v.reshape(-1).double() for v in sd.values()])
    return np.array([float(flat.sum()), float(flat.abs().sum()), float(flat.numel())])


def make_graph_attr(ref, max_neighbours=scene.MAX_NBRS):
    return ref.collate.GraphBoundedMultimaterialControl(
        scene.CONN_R, scene.make_stats(), scene.CART_IDX, [scene.MAT_IDX], scene.CTRL_IDX, scene.make_bounds(),
        noise=None, max_neighbours=max_neighbours)


def golden_graphs(ref):
    for name, h in (("graph_dense", 0.005), ("graph_sparse", 0.012)):
        pos, _ = scene.lattice_positions(600, h, seed=3)
        s, r = ref.utils.get_connectivity(pos, scene.CONN_R, scene.MAX_NBRS)
        np.savez_compressed(os.path.join(GOLDEN_DIR, name + ".npz"), pos=pos.numpy(), conn_r=scene.CONN_R,
                            max_neighbours=scene.MAX_NBRS, senders=s.numpy().astype(np.int32),
                            receivers=r.numpy().astype(np.int32))
        print(name, "edges", len(s))


def golden_step(ref):
    n = 400
    obs, nxt = scene.make_state(n, h=0.005, seed=5, n_rigid=120)
    g = torch.Generator().manual_seed(11)
    obs[-1, :120, 5:8] = torch.randn((120, 3), generator=g) * 3e-4      # non-trivial control columns
    ga = make_graph_attr(ref)
    nodes, edge_attr, edge_index, tgt = ga.process_collate([(obs, nxt)])
    torch.manual_seed(0)
    model = ref.EncProcDecGNN(nodes.shape[1], edge_attr.shape[1], 3, 128, 2, 10)
    with torch.no_grad():
        pred = model.forward(nodes, edge_attr, edge_index)
    next_pos = ref.rollout.get_position_from_prediction(ga.stats, ga.cartesian_idx, pred, obs)
    np.savez_compressed(os.path.join(GOLDEN_DIR, "step.npz"), obs=obs.numpy(), tgt=nxt.numpy(), nodes=nodes.numpy(),
                        edge_attr=edge_attr.numpy(), edge_index=edge_index.numpy().astype(np.int32),
                        target=tgt.numpy(), pred=pred.numpy(), next_pos=next_pos.numpy(),
                        weights_checksum=weights_checksum(model.state_dict()), weights_seed=0)
    print("step: nodes", tuple(nodes.shape), "edges", edge_attr.shape[0], "|pred|", float(pred.abs().mean()))


def golden_rollout(ref):
    n, n_rigid, horizon = 400, 120, 8
    obs, nxt = scene.make_state(n, h=0.005, seed=7, n_rigid=n_rigid)
    ga = make_graph_attr(ref)
    torch.manual_seed(0)
    model = ref.EncProcDecGNN(25, 4, 3, 128, 2, 10)
    rot = np.deg2rad(180.0 + 2.0 * np.arange(horizon))
    ty = -2e-4 * np.arange(horizon)
    preds, states = [], []

    class Recorder:
        def forward(self, x, e, ei):
            with torch.no_grad():
                p = model.forward(x, e, ei)
            preds.append(p.clone().numpy())
            return p

    class Solver(ref.traj.CMAESolver):               # only the abstract hooks; the loop is the reference's
        def interpolate_trajectory(self, x):
            return list(rot), list(ty)

        def compute_loss(self, end_position, actions, cup_states=None, coffee_states=None, x=None):
            states.append(end_position.clone())
            return 0.0, None, None, None, None, None

    with contextlib.redirect_stdout(io.StringIO()):
        solver = Solver(Recorder(), ga, (obs.clone(), nxt.clone()), scene.RX_INIT_DEG, scene.TY_INIT, 0.05, 1.0,
                        1.0, 1.0, 1.0, 1.0, 1.0, "cpu", total_steps=horizon, traj_points=1)
    # capture the final window: cma_objective clones initial_state, so wrap clone bookkeeping via compute_loss only;
    # the full end window is re-derived below with the same reference calls for the fixture.
    final = {}
    orig = ref.traj.get_position_from_prediction

    def spy(stats, cart, pred, window):
        final["window"] = window          # the live state tensor the reference mutates in place
        return orig(stats, cart, pred, window)

    ref.traj.get_position_from_prediction = spy
    try:
        solver.cma_objective(np.zeros(2))
    finally:
        ref.traj.get_position_from_prediction = orig
    traj = torch.stack([solver.compute_particles_tmatrix(rot[i], ty[i]) for i in range(horizon)])
    np.savez_compressed(os.path.join(GOLDEN_DIR, "rollout.npz"), obs0=obs.numpy(), rot=rot, ty=ty,
                        ty_init=np.array(scene.TY_INIT), n_rigid=n_rigid, horizon=horizon,
                        traj=traj.numpy(), preds=np.stack(preds), final_obs=final["window"].numpy(),
                        end_position=states[0].numpy(),
                        weights_checksum=weights_checksum(model.state_dict()), weights_seed=0)
    print("rollout: steps", len(preds), "final obs", tuple(final["window"].shape))


def _bench_control_curve(steps):
    """The control curve bench.py rolls out: dataset/sample_traj.npy, rotation in radians and the translation
    relative to its first row (same as bench.control_curve(steps, rank=0))."""
    st = np.load(os.path.join(ref_shims.REFERENCE_ROOT, "dataset", "sample_traj.npy"))
    reps = -(-(steps + 1) // st.shape[0])
    st = np.concatenate([st] * reps)[: steps + 1]
    return np.deg2rad(st[:steps, 0]), st[:steps, 1] - st[0, 1]


def golden_rollout_c2(ref, horizon=300, marks=(10, 50, 150, 300)):
    """SURVEY.md section 8 row D gate (4) at C2 size: the reference's own cma_objective loop, 2 000 particles
    (828 rigid), 300 free-running steps; the sand positions entering steps 10 / 50 / 150 and the end state are kept,
    together with the edge count of every step."""
    n, n_rigid = 2000, scene.N_RIGID
    obs, nxt = scene.make_state(n, seed=0, n_rigid=n_rigid)
    ga = make_graph_attr(ref)
    torch.manual_seed(0)
    model = ref.EncProcDecGNN(25, 4, 3, 128, 2, 10)
    rot, ty = _bench_control_curve(horizon)
    n_edges, out = [], {}

    class Recorder:
        def forward(self, x, e, ei):
            n_edges.append(int(ei.shape[1]))
            with torch.no_grad():
                return model.forward(x, e, ei)

    class Solver(ref.traj.CMAESolver):
        def interpolate_trajectory(self, x):
            return list(rot), list(ty)

        def compute_loss(self, end_position, actions, cup_states=None, coffee_states=None, x=None):
            for t in marks:
                out[t] = (end_position if t >= len(coffee_states) else coffee_states[t]).clone().numpy()
            return 0.0, None, None, None, None, None

    with contextlib.redirect_stdout(io.StringIO()):
        solver = Solver(Recorder(), ga, (obs.clone(), nxt.clone()), scene.RX_INIT_DEG, scene.TY_INIT, 0.05, 1.0,
                        1.0, 1.0, 1.0, 1.0, 1.0, "cpu", total_steps=horizon, traj_points=1)
    # full state windows entering steps 0 / 100 / 200 / 294 (control columns of that step already written): start
    # states of the CPU-baseline samples in bench.py, so that the timed CPU steps cover the horizon the GPU rolls out
    starts, windows, step_no = (0, 100, 200, 294), {}, [0]
    orig = ref.traj.get_position_from_prediction

    def spy(stats, cart, pred, window):
        if step_no[0] in starts:
            windows[step_no[0]] = window.clone().numpy()
        step_no[0] += 1
        return orig(stats, cart, pred, window)

    ref.traj.get_position_from_prediction = spy
    try:
        solver.cma_objective(np.zeros(2))
    finally:
        ref.traj.get_position_from_prediction = orig
    np.savez_compressed(os.path.join(GOLDEN_DIR, "rollout_c2.npz"), horizon=horizon, marks=np.array(marks),
                        n_rigid=n_rigid, rot=rot, ty=ty, n_edges=np.array(n_edges, dtype=np.int32),
                        weights_seed=0, starts=np.array(starts), **{f"sand_{t}": out[t] for t in marks},
                        **{f"obs_{t}": windows[t] for t in starts})
    print("rollout_c2: edges per step first/mean/last", n_edges[0], float(np.mean(n_edges)), n_edges[-1])


def golden_c3_forward(ref, n=100_000, stride=25):
    """One forward of the reference model on the C3 scene (100 000 particles, about 2 M edges); every `stride`-th
    node's normalised acceleration is kept -- bench.py checks its tensor-core forward against it before it reports
    a roofline number for that regime."""
    obs, nxt = scene.make_state(n, seed=0, n_rigid=0)
    ga = make_graph_attr(ref)
    nodes, edge_attr, edge_index, _ = ga.process_collate([(obs, nxt)])
    torch.manual_seed(0)
    model = ref.EncProcDecGNN(25, 4, 3, 128, 2, 10)
    with torch.no_grad():
        pred = model.forward(nodes, edge_attr, edge_index)
    idx = np.arange(0, n, stride)
    np.savez_compressed(os.path.join(GOLDEN_DIR, "c3_forward.npz"), n=n, n_edges=int(edge_index.shape[1]),
                        idx=idx.astype(np.int32), pred=pred.numpy()[idx], pred_norm=float(pred.norm()),
                        weights_seed=0)
    print("c3_forward: edges", int(edge_index.shape[1]), "|pred|", float(pred.abs().mean()))


def _make_solver(ref, cls, state, total_steps, traj_points, max_rot=3.0, max_ty=0.002, scale_rot=0.05):
    """A reference CMA-ES solver with the README defaults (README.md:221-224, examples/optimise_traj.py:250-251);
    the model is never called by the functions recorded here."""
    ga = make_graph_attr(ref)
    with contextlib.redirect_stdout(io.StringIO()):
        return cls(None, ga, state, scene.RX_INIT_DEG, scene.TY_INIT, scale_rot, 1.0, 1.0, 1.0, 1.0, 1.0, 1.0, "cpu",
                   cma_var=1.5, cma_rand=1234, max_rot=max_rot, max_ty=max_ty, total_steps=total_steps,
                   traj_points=traj_points)


def golden_traj_param(ref):
    """Row N2 (+ the regulariser half of N1): TrajectoryCMAsolver / InterpolatedCMAsolver parametrisations
    (gnn_manip/utils/traj_utils.py:199-228, 296-303, 393-416) and their loss terms (:230-246, 263-285, 353-373,
    375-391, 418-452), evaluated by the unmodified reference classes on dataset/sample_traj.npy."""
    st = np.load(os.path.join(ref_shims.REFERENCE_ROOT, "dataset", "sample_traj.npy"))
    obs, nxt = scene.make_state(400, h=0.005, seed=7, n_rigid=120)
    state = (obs.clone(), nxt.clone())
    out = {}

    def terms(solver, X, with_x):
        rows = []
        for x in X:
            rot, ty = solver.interpolate_trajectory(x)
            actions = np.zeros((solver.horizon, 2))
            actions[:, 0] = rot
            actions[:, 1] = ty
            vel, acc = solver.compute_vel_acc(actions)
            row = [solver.compute_vel_loss(vel), solver.compute_acc_loss(acc)]
            row.append(solver.compute_vel_noninterp(x) if with_x else solver.compute_boundaries_penalty(actions))
            rows.append(row)
        return np.array(rows, dtype=np.float64)

    # increments parametrisation (horizon 300, 598 parameters)
    inc = _make_solver(ref, ref.traj.TrajectoryCMAsolver, state, 300, 10)
    inc.set_sample_traj(st)
    x0 = np.concatenate((inc.sample_traj[:, 0], inc.sample_traj[:, 1]))
    rng = np.random.RandomState(1234)
    X = np.stack([x0, x0 + 0.01 * rng.randn(x0.size), x0 + 1.5 * rng.randn(x0.size)])
    curves = [inc.interpolate_trajectory(x) for x in X]
    out.update(inc_x=X, inc_rot=np.array([c[0] for c in curves]), inc_ty=np.array([c[1] for c in curves]),
               inc_terms=terms(inc, X, False), inc_scale_ty=inc.scale_ty)

    # pchip parametrisation, the C4 configuration of SURVEY.md section 0 F6: horizon 150, 3 knots, 6 parameters
    p150 = _make_solver(ref, ref.traj.InterpolatedCMAsolver, state, 150, 50)
    p150.set_sample_traj(st[:151])
    x0 = np.concatenate((p150.sample_traj[:, 0], p150.sample_traj[:, 1]))
    rng = np.random.RandomState(1234)
    X = np.concatenate((x0[None], x0[None] + 1.5 * rng.randn(15, x0.size)))
    curves = [p150.interpolate_trajectory(x) for x in X]
    rigid, actions = p150.get_rigid_body_trajectory_from_diff(X[3])
    out.update(p150_x=X, p150_rot=np.array([c[0] for c in curves]), p150_ty=np.array([c[1] for c in curves]),
               p150_terms=terms(p150, X, True), p150_rigid_x3=rigid.numpy(), p150_obs0=obs.numpy(),
               p150_ineq=np.array([p150.ineq_constraint(x) for x in X]))

    # pchip parametrisation at horizon 300 with 30 knots (60 parameters): non-monotone data, zero slopes
    p300 = _make_solver(ref, ref.traj.InterpolatedCMAsolver, state, 300, 10)
    p300.set_sample_traj(st)
    x0 = np.concatenate((p300.sample_traj[:, 0], p300.sample_traj[:, 1]))
    rng = np.random.RandomState(7)
    X = np.concatenate((x0[None], x0[None] + 0.02 * rng.randn(5, x0.size)))
    X[5, 3:6] = X[5, 2]                                    # flat segments: pchip slopes of exactly zero
    curves = [p300.interpolate_trajectory(x) for x in X]
    out.update(p300_x=X, p300_rot=np.array([c[0] for c in curves]), p300_ty=np.array([c[1] for c in curves]),
               p300_terms=terms(p300, X, True))
    np.savez_compressed(os.path.join(GOLDEN_DIR, "traj_param.npz"), **out)
    print("traj_param:", {k: np.asarray(v).shape for k, v in out.items()})


def golden_blocks(ref):
    tg = ref.torch_graphnet
    nodes = torch.tensor([[10.0, 20.0, 30.0], [11.0, 21.0, 31.0], [12.0, 22.0, 32.0], [13.0, 23.0, 33.0],
                          [14.0, 24.0, 34.0]])
    edges = torch.tensor([[100.0, 200.0], [101.0, 201.0], [102.0, 202.0], [103.0, 203.0], [104.0, 204.0],
                          [105.0, 205.0]])
    senders = torch.tensor([0, 1, 1, 2, 3, 4])
    receivers = torch.tensor([1, 2, 3, 0, 4, 3])
    ei = torch.stack((senders, receivers))
    torch.manual_seed(123)
    lin_e = torch.nn.Linear(2 + 3 + 3, 4)
    lin_n = torch.nn.Linear(4 + 3, 5)
    em = tg.EdgeModel(torch.nn.Sequential(lin_e))
    nm = tg.NodeModel(torch.nn.Sequential(lin_n))
    inet = tg.InteractionNetwork(torch.nn.Sequential(lin_e), torch.nn.Sequential(lin_n))
    with torch.no_grad():
        e_out = em(nodes, edges, ei)
        n_out = nm(nodes, e_out, ei)
        i_nodes, i_edges, _ = inet(nodes, edges, ei)
        agg_recv = tg.utils.gn_utils.received_edges_to_node_aggregator(nodes, edges, receivers, "sum") \
            if hasattr(tg, "utils") else torch.zeros(5, 2).index_add_(0, receivers, edges)
    np.savez_compressed(os.path.join(GOLDEN_DIR, "blocks.npz"), nodes=nodes.numpy(), edges=edges.numpy(),
                        senders=senders.numpy().astype(np.int32), receivers=receivers.numpy().astype(np.int32),
                        w_edge=lin_e.weight.detach().numpy(), b_edge=lin_e.bias.detach().numpy(),
                        w_node=lin_n.weight.detach().numpy(), b_node=lin_n.bias.detach().numpy(),
                        edge_out=e_out.numpy(), node_out=n_out.numpy(), inet_nodes=i_nodes.numpy(),
                        inet_edges=i_edges.numpy(), agg_recv=agg_recv.numpy())
    print("blocks: edge_out", tuple(e_out.shape), "node_out", tuple(n_out.shape))


def main(argv=None):
    """python -m oracle.make_golden [name ...]: all fixtures, or only the named ones (graphs, step, rollout, blocks,
    rollout_c2, c3_forward, traj_param, sample_traj)."""
    import sys
    names = list(sys.argv[1:] if argv is None else argv)
    os.makedirs(GOLDEN_DIR, exist_ok=True)
    ref = ref_shims.load()
    small = {"graphs": golden_graphs, "step": golden_step, "rollout": golden_rollout, "blocks": golden_blocks,
             "traj_param": golden_traj_param}
    big = {"rollout_c2": golden_rollout_c2, "c3_forward": golden_c3_forward}
    for name, fn in small.items():
        if not names or name in names:
            torch.set_num_threads(1)      # fixed reduction order inside MKL for reproducible fixtures
            fn(ref)
    for name, fn in big.items():          # minutes of CPU each: all host threads
        if not names or name in names:
            torch.set_num_threads(os.cpu_count() or 1)
            fn(ref)
    if not names or "sample_traj" in names:
        # the CMA-ES warm-start control curve (301 x 2 float64 data file, not source): input fixture for bench.py / tests
        import shutil
        shutil.copy(os.path.join(ref_shims.REFERENCE_ROOT, "dataset", "sample_traj.npy"),
                    os.path.join(GOLDEN_DIR, "sample_traj.npy"))


if __name__ == "__main__":
    main()
