"""Synthetic scenes, stats and weights of SURVEY.md section 8 row D.  TEST INFRASTRUCTURE ONLY.
The input generators live in /scenes.py (shared with bench.py); this module adds the oracle-model weights."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from scenes import *  # noqa: F401,F403,E402
from scenes import (CART_IDX, CONN_R, CTRL_IDX, DATA_DIM, K_STEPS, MAT_IDX, MAX_NBRS, N_RIGID,  # noqa: F401,E402
                    RX_INIT_DEG, TY_INIT, lattice_positions, make_bounds, make_obs, make_state, make_stats,
                    sample_traj)


def init_weights_state_dict(node_dim=25, edge_dim=4, out_dim=3, hidden=128, m_steps=10, seed=0):
    """Random-init state dict with the reference layout, from the oracle's own model class."""
    from .restate import EncProcDecGNN
    torch.manual_seed(seed)
    m = EncProcDecGNN(node_dim, edge_dim, out_dim, hidden, 2, m_steps)
    return m.state_dict()
