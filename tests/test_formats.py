"""CPU suite for row N4: the on-disk formats (metadata.json, particle CSVs, sim_data.csv, rollout CSV) round-trip and,
where /root/reference is present, read back identically through the reference's own loaders."""
import json
import os

import numpy as np
import pytest
import torch

from oracle import ref_shims, scene


def write_dataset(root, n_sims=2, T=9, n=40, seed=0):
    from gnn_manip_b200 import coffee_dataset as cd
    rng = np.random.RandomState(seed)
    os.makedirs(os.path.join(root, "train"), exist_ok=True)
    sims = []
    for s in range(n_sims):
        p0 = 0.4 + 0.2 * rng.rand(n, 3)
        v = 1e-3 * rng.randn(n, 3)
        data = np.zeros((T, n, 5))
        data[:, :, 0] = np.arange(n)
        data[:, : n // 4, 1] = 1.0
        for t in range(T):
            data[t, :, 2:5] = p0 + t * v + 1e-5 * rng.randn(n, 3)
        sims.append(data)
        cd.write_particles_csv(os.path.join(root, "train", f"particles_{s + 1:06d}.csv"), data)
    np.savetxt(os.path.join(root, "train", "sim_data.csv"), np.array([[s + 1, 0] for s in range(n_sims)]),
               delimiter=",", fmt="%d")
    meta = cd.write_metadata(os.path.join(root, "metadata.json"), sims, [2, 3, 4], [5, 6, 7], 1,
                             [[0.1, 0.9], [0.1, 0.9], [0.1, 0.9]])
    return sims, meta


def test_metadata_and_particles_round_trip(tmp_path):
    from gnn_manip_b200 import coffee_dataset as cd
    root = str(tmp_path) + "/"
    sims, meta = write_dataset(root)
    data_dim, T, cart, ctrl, mat, bounds, stats = cd.read_metadata(root + "metadata.json")
    assert (data_dim, T, cart, ctrl, mat) == (5, 9, [2, 3, 4], [5, 6, 7], 1)
    assert torch.equal(bounds["lower_bounds"], torch.tensor([0.1, 0.1, 0.1]))
    assert torch.equal(bounds["upper_bounds"], torch.tensor([0.9, 0.9, 0.9]))
    vel = np.concatenate([np.diff(s[:, :, 2:5], axis=0).reshape(-1, 3) for s in sims])
    assert np.allclose(stats["velocity_mean"].numpy(), vel.mean(0)) and np.allclose(stats["velocity_std"].numpy(), vel.std(0))
    assert set(json.load(open(root + "metadata.json"))) >= {"cartesian_idx", "control_idx", "material_id", "bounds",
                                                            "sequence_length", "dim", "data_dim", "vel_mean", "vel_std",
                                                            "acc_mean", "acc_std"}
    back = cd.read_particles_csv(root + "train/particles_000002.csv", T, data_dim)
    assert back.shape == sims[1].shape and np.allclose(back, sims[1], rtol=1e-8)


def test_windows_and_datasets(tmp_path):
    from gnn_manip_b200 import coffee_dataset as cd
    root = str(tmp_path) + "/"
    sims, _ = write_dataset(root)
    ds = cd.CoffeeTestDataset(root, k=6, conn_r=0.015, split="train", use_control=True, sim_id=2)
    assert len(ds) == 9 - 6
    obs, nxt = ds[1]
    assert obs.shape == (6, 40, 8) and nxt.shape == (40, 3) and obs.dtype == torch.float32
    assert np.allclose(obs[:, :, :5].numpy(), sims[1][1:7].astype(np.float32))
    assert np.allclose(nxt.numpy(), sims[1][7, :, 2:5].astype(np.float32))
    rigid = obs[-1, :, 1] == 1
    assert torch.equal(obs[:, ~rigid, 5:8], torch.zeros_like(obs[:, ~rigid, 5:8]))
    assert torch.equal(obs[-1, rigid, 5:8], nxt[rigid] - obs[-1, rigid, 2:5])
    assert type(ds.graph_attr).__name__ == "GraphBoundedMultimaterialControl" and ds.graph_attr.max_neighbours == 20
    plain = cd.CoffeeTestDataset(root, k=6, conn_r=0.015, split="train", use_control=False, sim_id=1)
    assert plain[0][0].shape == (6, 40, 5) and type(plain.graph_attr).__name__ == "GraphBoundedMultimaterial"
    full = cd.CoffeeDataset(root, k=6, conn_r=0.015, split="train", use_control=True)
    assert len(full) == 2 * 3 and torch.equal(full[4][0], ds[1][0])


def test_rollout_csv_format(tmp_path):
    from gnn_manip_b200 import coffee_dataset as cd
    pred = np.zeros((3, 4, 8))
    pred[:, :, 0] = np.arange(4)
    pred[:, 2:, 1] = 1
    pred[:, :, 2:5] = np.linspace(0.1, 0.9, 36).reshape(3, 4, 3)
    path = cd.save_as_csv(pred, str(tmp_path / "out"))
    lines = open(path).read().strip().split("\n")
    assert os.path.basename(path) == "positions.csv" and len(lines) == 12
    assert lines[0] == "0,0,0.100000,0.122857,0.145714" and lines[-1].startswith("3,1,")


@pytest.mark.skipif(not ref_shims.reference_available(), reason="/root/reference is not present")
def test_formats_against_reference_loaders(tmp_path):
    import contextlib
    import io
    ref_shims.load()
    with contextlib.redirect_stdout(io.StringIO()):
        from gnn_manip.utils import coffee_dataset as ref_cd
    from gnn_manip_b200 import coffee_dataset as cd
    root = str(tmp_path) + "/"
    write_dataset(root)
    a, b = ref_cd.read_metadata(root + "metadata.json"), cd.read_metadata(root + "metadata.json")
    assert a[:5] == b[:5]
    for da, db in zip(a[5:], b[5:]):
        assert da.keys() == db.keys() and all(torch.equal(da[k], db[k]) for k in da)
    ra = ref_cd.CoffeeTestDataset(root, 6, 0.015, split="train", use_control=True, sim_id=1)
    rb = cd.CoffeeTestDataset(root, 6, 0.015, split="train", use_control=True, sim_id=1)
    assert len(ra) == len(rb)
    for i in range(len(ra)):
        assert torch.equal(ra[i][0], rb[i][0]) and torch.equal(ra[i][1], rb[i][1])
