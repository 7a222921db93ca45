"""CPU suite: the oracle restatement (oracle/) against the golden vectors recorded from the unmodified reference
(tests/golden, written by oracle/make_golden.py) and against scikit-learn's KDTree (the reference's third-party
graph builder).  Bit-exact for the graph, the features and the integrator; 2e-6 relative for fp32 MLP outputs
(same torch ops, possibly different MKL blocking)."""
import numpy as np
import pytest
import torch

from conftest import load_golden, rel_l2, weights_checksum
from oracle import scene


def _features(restate, connectivity=None):
    return restate.GraphFeatures(scene.CONN_R, scene.make_stats(), scene.make_bounds(), scene.CART_IDX,
                                 scene.MAT_IDX, scene.CTRL_IDX, scene.MAX_NBRS,
                                 connectivity or restate.get_connectivity)


@pytest.mark.parametrize("name", ["graph_dense", "graph_sparse"])
def test_graph_restatements_match_reference_golden(oracle_lib, name):
    g = load_golden(name)
    pos = torch.from_numpy(g["pos"])
    for fn in (oracle_lib.get_connectivity, oracle_lib.get_connectivity_bruteforce):
        s, r = fn(pos, float(g["conn_r"]), int(g["max_neighbours"]))
        assert np.array_equal(s.numpy(), g["senders"].astype(np.int64))
        assert np.array_equal(r.numpy(), g["receivers"].astype(np.int64))


@pytest.mark.parametrize("n,h", [(2000, 0.005), (1500, 0.012), (37, 0.004)])
def test_graph_c_oracle_matches_sklearn(oracle_lib, n, h):
    pos, _ = scene.lattice_positions(n, h, seed=n)
    s0, r0 = oracle_lib.get_connectivity_sklearn(pos.numpy(), scene.CONN_R, scene.MAX_NBRS)
    s1, r1 = oracle_lib.get_connectivity(pos, scene.CONN_R, scene.MAX_NBRS)
    assert torch.equal(s0, s1) and torch.equal(r0, r1)


def test_graph_edge_cases(oracle_lib):
    one = torch.tensor([[0.5, 0.5, 0.5]])
    s, r = oracle_lib.get_connectivity(one, 0.015, 20)
    assert s.tolist() == [0] and r.tolist() == [0]                      # the self loop is an edge
    far = torch.tensor([[0.0, 0.0, 0.0], [1.0, 1.0, 1.0]])
    s, r = oracle_lib.get_connectivity(far, 0.015, 20)
    assert s.tolist() == [0, 1] and r.tolist() == [0, 1]
    pos, _ = scene.lattice_positions(300, 0.005, seed=1)
    s, r = oracle_lib.get_connectivity(pos, 0.015, 1)                    # cap 1 keeps only the self loop
    assert torch.equal(s, r) and len(s) == 300


def test_features_and_step_match_reference_golden(oracle_lib):
    g = load_golden("step")
    obs, tgt = torch.from_numpy(g["obs"]), torch.from_numpy(g["tgt"])
    feats = _features(oracle_lib)
    nodes, edge_attr, edge_index, target = feats.process_collate([(obs, tgt)])
    assert np.array_equal(edge_index.numpy(), g["edge_index"].astype(np.int64))
    assert np.array_equal(nodes.numpy(), g["nodes"])                    # bit-exact
    assert np.array_equal(edge_attr.numpy(), g["edge_attr"])
    assert np.array_equal(target.numpy(), g["target"])
    torch.manual_seed(int(g["weights_seed"]))
    model = oracle_lib.EncProcDecGNN(25, 4, 3, 128, 2, 10)
    assert np.allclose(weights_checksum(model.state_dict()), g["weights_checksum"], rtol=1e-12)
    with torch.no_grad():
        pred = model(nodes, edge_attr, edge_index)
    assert rel_l2(pred, g["pred"]) < 2e-6
    nxt = oracle_lib.get_position_from_prediction(feats.stats, scene.CART_IDX, torch.from_numpy(g["pred"]), obs)
    assert np.array_equal(nxt.numpy(), g["next_pos"])


def test_state_dict_layout_matches_reference(oracle_lib):
    sd = oracle_lib.EncProcDecGNN(25, 4, 3, 128, 2, 10).state_dict()
    assert len(sd) == 342
    assert sd["processor.3.phi_edge.0.weight"].shape == (128, 384)
    assert sd["processor.3.edge_model.phi_edge.0.weight"].data_ptr() == sd["processor.3.phi_edge.0.weight"].data_ptr()
    assert sd["processor.9.node_model.phi_node.0.weight"].shape == (128, 256)
    assert sd["decoder.4.weight"].shape == (3, 128)
    unique = {v.data_ptr(): v.numel() for v in sd.values()}
    assert len(unique) == 182 and sum(unique.values()) == 1591299


def test_rollout_matches_reference_golden(oracle_lib):
    g = load_golden("rollout")
    obs0 = torch.from_numpy(g["obs0"])
    feats = _features(oracle_lib)
    torch.manual_seed(int(g["weights_seed"]))
    model = oracle_lib.EncProcDecGNN(25, 4, 3, 128, 2, 10)
    rigid = obs0[-1][obs0[-1][:, scene.MAT_IDX] == 1][:, scene.CART_IDX]
    traj = oracle_lib.rigid_body_trajectory(g["rot"], g["ty"], int(g["horizon"]), list(g["ty_init"]), rigid)
    assert np.array_equal(traj.numpy(), g["traj"])                       # rigid transform bit-exact
    preds = []
    final = oracle_lib.rollout(model, feats, (obs0, None), traj, feats.stats, int(g["horizon"]),
                               record=lambda i, d: preds.append(d["pred"].clone()))
    # free-running over 8 steps: fp32 MLP rounding may differ at 1e-6, positions move by acc_std * 1e-6
    assert rel_l2(torch.stack(preds), g["preds"]) < 1e-5
    assert np.abs(final.numpy() - g["final_obs"]).max() < 1e-7
    coffee = final[-1][final[-1][:, scene.MAT_IDX] == 0][:, scene.CART_IDX]
    assert np.abs(coffee.numpy() - g["end_position"]).max() < 1e-7


def test_blocks_known_answer(oracle_lib):
    """The 5-node / 6-edge fixture of deps/torch-graphnet/test/gnn_test.py:81-133, restated with numpy
    (concat + W x + b + segment sum) and checked against the reference's recorded outputs."""
    g = load_golden("blocks")
    nodes, edges = g["nodes"], g["edges"]
    s, r = g["senders"], g["receivers"]
    e_in = np.concatenate((edges, nodes[r], nodes[s]), axis=1)            # [edges, receivers, senders]
    e_out = e_in @ g["w_edge"].T + g["b_edge"]
    assert np.allclose(e_out, g["edge_out"], rtol=1e-5, atol=1e-3)
    agg = np.zeros((5, e_out.shape[1]), np.float32)
    np.add.at(agg, r, g["edge_out"])
    n_out = np.concatenate((agg, nodes), axis=1) @ g["w_node"].T + g["b_node"]
    assert np.allclose(n_out, g["node_out"], rtol=1e-5, atol=1e-3)
    assert np.allclose(g["inet_edges"], g["edge_out"]) and np.allclose(g["inet_nodes"], g["node_out"])
    agg_e = np.zeros((5, 2), np.float32)
    np.add.at(agg_e, r, edges)
    assert np.array_equal(agg_e, g["agg_recv"])
    # the oracle's torch classes reproduce the same numbers
    em = oracle_lib.EdgeModel(torch.nn.Sequential(torch.nn.Linear(8, 4)))
    em.phi_edge[0].weight.data = torch.from_numpy(g["w_edge"])
    em.phi_edge[0].bias.data = torch.from_numpy(g["b_edge"])
    ei = torch.from_numpy(np.stack((s, r)).astype(np.int64))
    with torch.no_grad():
        out = em(torch.from_numpy(nodes), torch.from_numpy(edges), ei)
    assert np.allclose(out.numpy(), g["edge_out"], rtol=1e-6, atol=1e-4)
