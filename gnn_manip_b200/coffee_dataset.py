"""On-disk formats of gnn-manip, so real datasets and rollouts flow through unchanged:

  metadata.json           gnn_manip/utils/coffee_dataset.py:18-43 (read_metadata), written by
                          simulation/generate_metadata.py:33-45
  particles_XXXXXX.csv    headerless [T * N, data_dim] rows, time-major (coffee_dataset.py:78); columns
                          [id, material, x, y, z] (scripts/render_dyn_blender.py:146-150)
  sim_data.csv            headerless (sim_id, _) rows listing the simulations of a split (coffee_dataset.py:51-52)
  positions.csv           rollout export `id,mat,x,y,z` with format %d,%d,%f,%f,%f (scripts/render_dyn.py:54-67)

Pure host I/O: no kernel work.  The windowed samples keep the reference's tensor layout -- obs [k, N, data_dim (+3
control columns)], next_pos [N, 3] -- which is what the rollout plan and the training path consume.
"""
import json
import os

import numpy as np
import torch


def read_metadata(metadata_file):
    """coffee_dataset.py:18-43 -> (data_dim, time_steps, cartesian_idx, control_idx, material_id, bounds, stats)."""
    with open(metadata_file) as f:
        metadata = json.load(f)
    data_dim = metadata["data_dim"]
    time_steps = metadata["sequence_length"]
    mb = torch.tensor(metadata["bounds"])
    bounds = {"upper_bounds": mb[:, 1], "lower_bounds": mb[:, 0]}
    stats = {"velocity_mean": torch.tensor(metadata["vel_mean"]), "velocity_std": torch.tensor(metadata["vel_std"]),
             "acceleration_mean": torch.tensor(metadata["acc_mean"]),
             "acceleration_std": torch.tensor(metadata["acc_std"])}
    return (data_dim, time_steps, metadata["cartesian_idx"], metadata["control_idx"], metadata["material_id"], bounds,
            stats)


def write_metadata(metadata_file, data, cartesian_idx, control_idx, material_id, bounds):
    """simulation/generate_metadata.py:20-45 for in-memory trajectories: data = list of [T, N, data_dim] arrays."""
    dim = len(cartesian_idx)
    vel_l, acc_l = [], []
    for d in data:
        vel = np.diff(np.asarray(d)[:, :, cartesian_idx], n=1, axis=0)
        acc = np.diff(vel, n=1, axis=0)
        vel_l.append(vel.reshape(-1, dim))
        acc_l.append(acc.reshape(-1, dim))
    vel, acc = np.concatenate(vel_l), np.concatenate(acc_l)
    metadata = {"cartesian_idx": list(cartesian_idx), "control_idx": list(control_idx), "material_id": material_id,
                "bounds": [list(map(float, b)) for b in bounds], "sequence_length": int(np.asarray(data[0]).shape[0]),
                "dim": dim, "data_dim": int(np.asarray(data[0]).shape[2]),
                "vel_mean": [float(v) for v in vel.mean(0)], "vel_std": [float(v) for v in vel.std(0)],
                "acc_mean": [float(v) for v in acc.mean(0)], "acc_std": [float(v) for v in acc.std(0)]}
    with open(metadata_file, "w") as f:
        json.dump(metadata, f)
    return metadata


def read_particles_csv(path, time_steps, data_dim):
    """coffee_dataset.py:78: headerless CSV -> float64 array [time_steps, N, data_dim]."""
    try:
        import pandas as pd
        flat = np.array(pd.read_csv(path, header=None))
    except ImportError:
        flat = np.loadtxt(path, delimiter=",", ndmin=2)
    return flat.reshape(time_steps, -1, data_dim)


def write_particles_csv(path, data):
    """Inverse of read_particles_csv (full precision)."""
    d = np.asarray(data)
    np.savetxt(path, d.reshape(-1, d.shape[-1]), delimiter=",", fmt="%.9g")


def save_as_csv(prediction_npy, output_dir):
    """scripts/render_dyn.py:54-67: rollout [T, N, >=5] -> <output_dir>/positions.csv with rows id,mat,x,y,z."""
    os.makedirs(output_dir, exist_ok=True)
    path = os.path.join(output_dir, "positions.csv")
    data = np.asarray(prediction_npy)[:, :, :5]
    dim = data.shape[-1]
    data = np.reshape(data, (-1, dim))
    np.savetxt(path, data, delimiter=",", fmt="%d,%d,%f,%f,%f" if dim == 5 else "%f")
    return path


def windows(data, k, cartesian_idx, material_id, use_control):
    """coffee_dataset.py:73-101 / 166-194: every (obs [k, N, D(+3)], next_pos [N, 3]) window of one simulation.
    Control columns (use_control): next_pos - pos for rigid particles (material == 1) at every frame of the window,
    zero for the rest (:89-97)."""
    data = np.asarray(data)
    pos = data[:, :, cartesian_idx]
    out = []
    for t in range(data.shape[0] - k):
        obs = torch.from_numpy(data[t: t + k]).float()
        nxt = torch.from_numpy(pos[t + k]).float()
        if use_control:
            no_control = obs[:, :, material_id] != 1
            ctrl = nxt - obs[:, :, cartesian_idx]
            ctrl[no_control] = 0
            obs = torch.cat((obs, ctrl.float()), dim=-1)
        out.append((obs, nxt))
    return out


class CoffeeTestDataset(torch.utils.data.Dataset):
    """coffee_dataset.py:136-221: raw (obs_seq, next_pos) samples of one simulation + the matching graph_attr."""

    def __init__(self, directory, k, conn_r, split="train", noise=None, max_neighbours=20, device=None,
                 use_control=False, sim_id=1):
        from .collate_utils import GraphBoundedMultimaterial, GraphBoundedMultimaterialControl
        self.dir, self.split = directory, split
        self.files = [f"{self.dir}{self.split}/particles_{sim_id:06d}.csv"]
        self.metadata_file = f"{self.dir}metadata.json"
        self.k, self.conn_r, self.max_neighbours = k, conn_r, max_neighbours
        self.use_control, self.device, self.noise = use_control, device, noise
        (self.data_dim, self.time_steps, self.cartesian_idx, self.control_idx, self.material_id, self.bounds,
         self.stats) = read_metadata(self.metadata_file)
        samples = []
        for f in self.files:
            samples += windows(read_particles_csv(f, self.time_steps, self.data_dim), k, self.cartesian_idx,
                               self.material_id, use_control)
        self.raw_samples = {"observations": [s[0] for s in samples], "next_positions": [s[1] for s in samples]}
        if use_control:
            self.graph_attr = GraphBoundedMultimaterialControl(conn_r, self.stats, self.cartesian_idx,
                                                               [self.material_id], self.control_idx, self.bounds,
                                                               noise=noise, max_neighbours=max_neighbours)
        else:
            self.graph_attr = GraphBoundedMultimaterial(conn_r, self.stats, self.cartesian_idx, [self.material_id],
                                                        self.bounds, noise=noise, max_neighbours=max_neighbours)

    def __len__(self):
        return len(self.raw_samples["observations"])

    def __getitem__(self, index):
        return self.raw_samples["observations"][index], self.raw_samples["next_positions"][index]


class CoffeeDataset(CoffeeTestDataset):
    """coffee_dataset.py:46-132: every simulation listed in <split>/sim_data.csv; samples stay raw windows -- the graph
    of a sample is built on the device when a batch is assembled (collate_utils.GraphAttributes.process_collate /
    the training step), not per item on the CPU as the reference's __getitem__ does."""

    def __init__(self, root, k, conn_r, split="train", noise=None, device=None, transform=None, use_control=False,
                 max_neighbours=20):
        assert split in ["train", "test"]
        sims = np.atleast_2d(np.loadtxt(f"{root}{split}/sim_data.csv", delimiter=","))
        first = int(sims[0, 0])
        super().__init__(root, k, conn_r, split=split, noise=noise, max_neighbours=max_neighbours, device=device,
                         use_control=use_control, sim_id=first)
        for sim_id in sims[1:, 0]:
            f = f"{root}{split}/particles_{int(sim_id):06d}.csv"
            self.files.append(f)
            for obs, nxt in windows(read_particles_csv(f, self.time_steps, self.data_dim), k, self.cartesian_idx,
                                    self.material_id, use_control):
                self.raw_samples["observations"].append(obs)
                self.raw_samples["next_positions"].append(nxt)
