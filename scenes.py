"""Synthetic scenes and stats of SURVEY.md section 8 row D (inputs only: no reference algorithm lives here).
Shared by bench.py, the tests and the oracle harness.

The reference's dataset (particle CSVs, metadata.json, trained_model.pth) is not in the
checkout (.MISSING_LARGE_BLOBS), so every input is synthetic.  Layout contracts followed:
  obs        [k=6, N, 8] f32, columns [id, material, x, y, z, cx, cy, cz]
             (scripts/render_dyn_blender.py:146-150, gnn_manip/utils/coffee_dataset.py:89-97)
  stats      velocity_mean/std, acceleration_mean/std, each [3] f32 (coffee_dataset.py:35-42)
  bounds     lower_bounds / upper_bounds [3] f32
"""
import numpy as np
import torch

K_STEPS = 6
DATA_DIM = 8
CART_IDX = [2, 3, 4]
MAT_IDX = 1
CTRL_IDX = [5, 6, 7]
CONN_R = 0.015
MAX_NBRS = 20
N_RIGID = 828           # the count in the comment at gnn_manip/utils/traj_utils.py:181
RX_INIT_DEG = 180.0     # examples/optimise_traj.py:250-251
TY_INIT = [0.5, 0.5, 0.4]


def make_stats():
    return {
        "velocity_mean": torch.tensor([0.0, -1e-4, 0.0], dtype=torch.float32),
        "velocity_std": torch.tensor([3e-4, 4e-4, 3e-4], dtype=torch.float32),
        "acceleration_mean": torch.tensor([0.0, -1e-6, 0.0], dtype=torch.float32),
        "acceleration_std": torch.tensor([5e-5, 6e-5, 5e-5], dtype=torch.float32),
    }


def make_bounds():
    return {
        "lower_bounds": torch.full((3,), 0.1, dtype=torch.float32),
        "upper_bounds": torch.full((3,), 0.9, dtype=torch.float32),
    }


def lattice_positions(n, h, seed, origin=None):
    """Jittered cubic lattice in general position (exact ties make the reference's order undefined)."""
    g = torch.Generator().manual_seed(seed)
    side = int(round(n ** (1.0 / 3.0))) + 1
    if origin is None:
        origin = 0.45 if n <= 4000 else 0.30
    ax = torch.arange(side, dtype=torch.float32)
    ii, jj, kk = torch.meshgrid(ax, ax, ax, indexing="ij")
    grid = torch.stack((ii.reshape(-1), jj.reshape(-1), kk.reshape(-1)), dim=1)[:n]
    jitter = (torch.rand((n, 3), generator=g, dtype=torch.float32) - 0.5) * (h / 2.0)
    return (origin + h * grid + jitter).float(), g


def make_obs(n, h=0.005, seed=0, n_rigid=N_RIGID, origin=None):
    """obs [6, N, 8] f32: k frames pos_t = p + t*v, v ~ N(0, (2e-4)^2); control columns zero."""
    p, g = lattice_positions(n, h, seed, origin)
    v = torch.randn((n, 3), generator=g, dtype=torch.float32) * 2e-4
    obs = torch.zeros((K_STEPS, n, DATA_DIM), dtype=torch.float32)
    obs[:, :, 0] = torch.arange(n, dtype=torch.float32)
    obs[:, : min(n_rigid, n), MAT_IDX] = 1.0
    for t in range(K_STEPS):
        obs[t, :, 2:5] = p + float(t) * v
    return obs


def make_state(n, h=0.005, seed=0, n_rigid=N_RIGID, origin=None):
    """(obs, next_pos) tuple as CoffeeTestDataset.__getitem__ returns (coffee_dataset.py:196-199)."""
    obs = make_obs(n, h, seed, n_rigid, origin)
    nxt = obs[-1, :, 2:5] + (obs[-1, :, 2:5] - obs[-2, :, 2:5])
    return obs, nxt.clone()


def sample_traj(rows=301):
    """Stand-in for dataset/sample_traj.npy when /root/reference is absent (GPU box): a smooth
    (rot_deg, ty) control curve of the same shape/range (first row [180, 0.5])."""
    t = np.linspace(0.0, 1.0, rows)
    rot = 180.0 + 45.0 * np.sin(2.0 * np.pi * t) * np.minimum(1.0, 4.0 * t)
    ty = 0.5 - 0.012 * np.sin(np.pi * t)
    return np.stack((rot, ty), axis=1)
