"""gnn_manip_b200 -- B200-native (sm_100a) implementation of the gnn-manip GraphNet rollout hot path.

Mirrors the reference's Python interface for that path (same names, arguments and error behaviour):
  gnn_manip.models.EncProcDecGNN                    -> gnn_manip_b200.EncProcDecGNN
  torch_graphnet.{EdgeModel, NodeModel, ...}        -> gnn_manip_b200.graph_networks
  gnn_manip.utils.utils / collate_utils / rollout_utils -> gnn_manip_b200.utils / collate_utils / rollout_utils
Everything computes through libgnn_manip_b200.so (C ABI: include/gnn_manip_b200.h); there is no CPU fallback.
"""
from . import _lib  # noqa: F401
from .epd_gnn import EncProcDecGNN  # noqa: F401
from .graph_networks import (EdgeModel, GraphIndependent, GraphNetwork, InteractionNetwork,  # noqa: F401
                             NodeModel)
from .collate_utils import (GraphAttributes, GraphBoundedMultimaterial,  # noqa: F401
                            GraphBoundedMultimaterialControl, GraphSimple)
from .rollout_utils import (Rollout, compute_rollout, get_position_from_prediction, perform_prediction,  # noqa: F401
                            rollout_final_state)
from .coffee_dataset import CoffeeDataset, CoffeeTestDataset, read_metadata, save_as_csv  # noqa: F401
from .traj_utils import (CMAESolver, InterpolatedCMAsolver, TrajectoryCMAsolver,  # noqa: F401
                         sinkhorn_divergence)
from .train_dyn import Batch, Trainer, average_gradients, make_batch  # noqa: F401
from .utils import (compute_acceleration, get_connectivity, get_edges_displacement,  # noqa: F401
                    get_nodes_vel, radius_graph, random_walk_noise)

__version__ = "0.1.0"
