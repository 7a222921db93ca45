"""GPU suite (-m gpu): compute_rollout with the reference contract (prediction [T, N, D]) over a dataset read from disk
(rows B + N4), against the oracle's loop."""
import types

import numpy as np
import pytest
import torch

from oracle import scene
from test_formats import write_dataset
from test_gpu_parity import seeded_models

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gm():
    import gnn_manip_b200
    assert torch.cuda.is_available()
    return gnn_manip_b200


def _oracle_feats(oracle_lib, ds):
    return oracle_lib.GraphFeatures(ds.conn_r, ds.stats, ds.bounds, ds.cartesian_idx, ds.material_id, ds.control_idx, 20)


def test_compute_rollout_cma_traj_mode(gm, oracle_lib, tmp_path):
    root = str(tmp_path) + "/"
    write_dataset(root, n_sims=1, T=9, n=300, seed=3)
    ds = gm.CoffeeTestDataset(root, k=6, conn_r=0.05, split="train", use_control=True, sim_id=1)
    ref, mine = seeded_models(gm, oracle_lib, seed=0)
    T = ds.time_steps
    actions = np.stack((np.deg2rad(180.0 + 0.7 * np.arange(T)), -1e-4 * np.arange(T)), axis=1)
    args = types.SimpleNamespace(cma_traj=actions, k_steps=6, device="cuda")
    pred = gm.compute_rollout(ds, mine, args)
    assert pred.shape == (T, 300, 8) and pred.dtype == np.float64
    obs0 = ds[0][0]
    feats = _oracle_feats(oracle_lib, ds)
    rigid = obs0[-1, :, 1] == 1
    traj = oracle_lib.rigid_body_trajectory(actions[:, 0], actions[:, 1], T, [0.5, 0.5, 0.4], obs0[-1, rigid][:, [2, 3, 4]])
    frames = []
    oracle_lib.rollout(ref, feats, (obs0, None), traj, feats.stats, T,
                       record=lambda i, d: frames.append(d["obs"][-1].clone()))
    want = torch.stack(frames).numpy()
    assert np.array_equal(pred[..., :2], want[..., :2])
    assert np.abs(pred - want).max() < 2e-6
    assert torch.equal(ds[0][0], obs0), "the dataset's stored sample must not be advanced"


def test_compute_rollout_ground_truth_mode(gm, oracle_lib, tmp_path):
    root = str(tmp_path) + "/"
    write_dataset(root, n_sims=1, T=10, n=200, seed=5)
    ds = gm.CoffeeTestDataset(root, k=6, conn_r=0.05, split="train", use_control=True, sim_id=1)
    ref, mine = seeded_models(gm, oracle_lib, seed=1)
    args = types.SimpleNamespace(cma_traj=None, k_steps=6, device="cuda")
    pred = gm.compute_rollout(ds, mine, args)
    steps = ds.time_steps - 6
    assert pred.shape == (steps, 200, 8)
    # the reference loop (rollout_utils.py:38-65) restated with the oracle's pieces
    feats = _oracle_feats(oracle_lib, ds)
    gt = torch.stack([ds[i][0][-1] for i in range(steps)])
    cur = ds[0][0].clone()
    rigid = cur[0, :, 1] == 1
    want = np.zeros_like(pred)
    with torch.no_grad():
        for i in range(steps):
            nr = cur[-1, rigid].clone()
            nr[:, 5:8] = gt[i, rigid][:, 5:8]
            cur[-1, rigid] = nr
            want[i] = cur[-1].numpy()
            nodes, ea, s, r, _ = feats.process(cur)
            acc = ref(nodes, ea, torch.stack((s, r)))
            nxt = oracle_lib.get_position_from_prediction(feats.stats, [2, 3, 4], acc, cur)
            cur[:-1] = cur[1:].clone()
            cur[-1][:, 2:5] = nxt
            nr[:, 2:5] = gt[i, rigid][:, 2:5]
            cur[-1, rigid] = nr
    assert np.abs(pred - want).max() < 2e-6


def test_save_rollout_csv_from_device_rollout(gm, oracle_lib, tmp_path):
    root = str(tmp_path) + "/"
    write_dataset(root, n_sims=1, T=8, n=120, seed=7)
    ds = gm.CoffeeTestDataset(root, k=6, conn_r=0.05, split="train", use_control=True, sim_id=1)
    _, mine = seeded_models(gm, oracle_lib, seed=0)
    T = ds.time_steps
    actions = np.stack((np.full(T, np.pi), np.zeros(T)), axis=1)
    pred = gm.compute_rollout(ds, mine, types.SimpleNamespace(cma_traj=actions, k_steps=6, device="cuda"))
    path = gm.save_as_csv(pred, str(tmp_path / "render"))
    back = np.loadtxt(path, delimiter=",").reshape(T, 120, 5)
    assert np.array_equal(back[..., :2], pred[..., :2]) and np.abs(back[..., 2:] - pred[..., 2:5]).max() < 1e-6
