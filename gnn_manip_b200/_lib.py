"""ctypes binding of libgnn_manip_b200.so (C ABI declared in include/gnn_manip_b200.h).

There is no CPU fallback: importing works without a GPU (so the symbol table can be checked), but every
compute entry point needs the CUDA library AND a CUDA device, and raises otherwise.
"""
import ctypes
import os
from ctypes import POINTER, c_char_p, c_double, c_float, c_int32, c_int64, c_longlong, c_size_t, c_void_p

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
# GNN_MANIP_B200_LIB selects a debug variant of the library (csrc/Makefile VARIANT=...), e.g. the phase-stamp build
LIB_PATH = os.environ.get("GNN_MANIP_B200_LIB") or os.path.join(_HERE, "libgnn_manip_b200.so")

GM_IMPL_AUTO, GM_IMPL_SIMT, GM_IMPL_TC = 0, 1, 2


class GmError(RuntimeError):
    """A C-ABI call returned a non-zero gm_status."""


class gm_phys(ctypes.Structure):
    _fields_ = [
        ("cartesian_idx", c_int32 * 3), ("material_idx", c_int32), ("control_idx", c_int32 * 3),
        ("use_control", c_int32),
        ("velocity_mean", c_float * 3), ("velocity_std", c_float * 3),
        ("acceleration_mean", c_float * 3), ("acceleration_std", c_float * 3),
        ("lower_bounds", c_float * 3), ("upper_bounds", c_float * 3),
        ("conn_r", c_double),
    ]


class gm_rollout_desc(ctypes.Structure):
    _fields_ = [
        ("model", c_void_p), ("impl", c_int32), ("phys", gm_phys),
        ("k", c_int32), ("D", c_int32), ("n_per_graph", c_int32), ("n_graphs", c_int32),
        ("max_neighbours", c_int32), ("n_rigid", c_int32), ("rigid_idx", c_void_p),
        ("traj_steps", c_int32), ("traj", c_void_p), ("rot", c_void_p), ("ty", c_void_p),
        ("rigid_init", c_void_p), ("ty_init", c_float * 3), ("obs", c_void_p), ("pred", c_void_p),
        ("use_cuda_graph", c_int32),
    ]


class gm_traj_param(ctypes.Structure):
    _fields_ = [
        ("kind", c_int32), ("horizon", c_int32), ("n_points", c_int32), ("knot_stride", c_int32),
        ("rx_init", c_double), ("scale_rot", c_double), ("scale_ty", c_double),
        ("max_rot", c_double), ("max_ty", c_double), ("rotation_limit", c_double), ("bound_penalty", c_double),
        ("vel_div", c_double * 2), ("acc_div", c_double * 2),
    ]


GM_TRAJ_INCREMENTS, GM_TRAJ_PCHIP = 0, 1


class gm_train_desc(ctypes.Structure):
    _fields_ = [("node_dim", c_int32), ("edge_dim", c_int32), ("out_dim", c_int32), ("hidden", c_int32),
                ("num_layers", c_int32), ("m_steps", c_int32), ("max_nodes", c_int32), ("max_edges", c_int32),
                ("params", c_void_p), ("grads", c_void_p)]


GRAD_READY_FN = ctypes.CFUNCTYPE(None, c_void_p, c_int32, c_size_t, c_size_t)


class gm_pack_job(ctypes.Structure):
    _fields_ = [("w", c_void_p), ("ld", c_int32), ("col0", c_int32), ("transpose", c_int32), ("chunk", c_void_p)]

# name -> (restype, argtypes); every symbol of include/gnn_manip_b200.h
SIGNATURES = {
    "gm_version": (ctypes.c_int, []),
    "gm_last_error": (c_char_p, []),
    "gm_launch_count": (c_longlong, []),
    "gm_radius_graph_workspace_bytes": (c_size_t, [c_int32, c_int32, c_int32]),
    "gm_radius_graph": (ctypes.c_int, [c_void_p, c_int32, c_int32, c_int32, c_double, c_int32] + [c_void_p] * 9 +
                        [c_size_t, c_void_p]),
    "gm_edge_index_i64": (ctypes.c_int, [c_void_p, c_void_p, c_void_p, c_int32, c_int32, c_int64, c_void_p, c_void_p]),
    "gm_recv_csr_workspace_bytes": (c_size_t, [c_int32, c_int64]),
    "gm_recv_csr": (ctypes.c_int, [c_void_p, c_void_p, c_int64, c_int32] + [c_void_p] * 5 + [c_size_t, c_void_p]),
    "gm_node_dim": (c_int32, [c_int32, c_int32]),
    "gm_node_features": (ctypes.c_int, [c_void_p, c_int32, c_int32, c_int32, POINTER(gm_phys), c_void_p, c_void_p]),
    "gm_edge_features": (ctypes.c_int, [c_void_p, c_int32, c_void_p, c_void_p, c_void_p, c_int32, c_double, c_void_p,
                                        c_void_p]),
    "gm_model_num_params": (c_size_t, [c_int32] * 6),
    "gm_model_blob_bytes": (c_size_t, [c_int32] * 6),
    "gm_model_create": (ctypes.c_int, [POINTER(c_void_p), c_void_p, c_size_t] + [c_int32] * 6 +
                        [c_void_p, c_size_t, c_void_p]),
    "gm_model_destroy": (None, [c_void_p]),
    "gm_model_supports_tc": (ctypes.c_int, [c_void_p]),
    "gm_forward_workspace_bytes": (c_size_t, [c_void_p, c_int32, c_int32, c_int32]),
    "gm_model_forward": (ctypes.c_int, [c_void_p, c_int32, c_void_p, c_void_p, c_void_p, c_int32, c_double, c_int32,
                                        c_int32] + [c_void_p] * 7 + [c_size_t, c_void_p]),
    "gm_encode": (ctypes.c_int, [c_void_p, c_int32, c_void_p, c_void_p, c_void_p, c_int32, c_double, c_int32, c_int32] +
                  [c_void_p] * 6 + [c_size_t, c_void_p]),
    "gm_mp_step": (ctypes.c_int, [c_void_p, c_int32, c_int32, c_int32, c_int32] + [c_void_p] * 5 + [c_size_t, c_void_p]),
    "gm_decode": (ctypes.c_int, [c_void_p, c_int32, c_int32, c_int32, c_void_p, c_void_p, c_size_t, c_void_p]),
    "gm_model_forward_profile": (ctypes.c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int32, c_double, c_int32,
                                                c_int32] + [c_void_p] * 7 + [c_size_t, c_void_p, POINTER(c_float),
                                                                             POINTER(c_int32)]),
    "gm_mlp_forward": (ctypes.c_int, [c_int32, POINTER(c_void_p), POINTER(c_void_p), POINTER(c_int32), c_void_p,
                                      c_void_p, c_int32, POINTER(c_void_p), POINTER(c_void_p), POINTER(c_int32),
                                      c_int64, c_void_p, c_void_p]),
    "gm_segment_sum": (ctypes.c_int, [c_void_p, c_int32, c_void_p, c_void_p, c_int32, c_void_p, c_void_p]),
    "gm_next_position": (ctypes.c_int, [c_void_p, c_int32, c_int32, c_int32, POINTER(gm_phys), c_void_p, c_void_p,
                                        c_void_p]),
    "gm_rollout_workspace_bytes": (c_size_t, [POINTER(gm_rollout_desc)]),
    "gm_rollout_create": (ctypes.c_int, [POINTER(c_void_p), POINTER(gm_rollout_desc), c_void_p, c_size_t, c_void_p]),
    "gm_rollout_run": (ctypes.c_int, [c_void_p, c_int32, c_void_p]),
    "gm_rollout_reset": (ctypes.c_int, [c_void_p, c_int32, c_void_p]),
    "gm_rollout_step_profile": (ctypes.c_int, [c_void_p, c_void_p, POINTER(c_float), POINTER(c_int32)]),
    "gm_rollout_graph": (ctypes.c_int, [c_void_p] + [POINTER(c_void_p)] * 5),
    "gm_rollout_kernel_launches_per_step": (c_int64, [c_void_p]),
    "gm_rollout_destroy": (None, [c_void_p]),
    "gm_traj_interpolate": (ctypes.c_int, [POINTER(gm_traj_param), c_void_p, c_int32, c_void_p, c_void_p, c_void_p,
                                           c_void_p]),
    "gm_rigid_trajectory": (ctypes.c_int, [c_void_p, c_void_p, c_int32, c_int32, c_void_p, c_int32, POINTER(c_float),
                                           c_void_p, c_void_p]),
    "gm_sinkhorn_workspace_bytes": (c_size_t, [c_int32, c_int32, c_int32]),
    "gm_sinkhorn_divergence": (ctypes.c_int, [c_void_p, c_int32, c_int64, c_void_p, c_int32, c_void_p, c_int32, c_int32,
                                              c_double, c_double, c_double, c_void_p, c_void_p, c_size_t, c_void_p]),
    "gm_train_pack": (ctypes.c_int, [c_void_p, c_int32, c_void_p]),
    "gm_train_gemm": (ctypes.c_int, [c_void_p] * 9 + [c_int32, c_int32, c_int32, c_void_p, c_void_p, c_void_p]),
    "gm_train_wgrad_partial_bytes": (c_size_t, []),
    "gm_train_wgrad": (ctypes.c_int, [c_void_p, c_void_p, c_int32, c_void_p, c_void_p, c_int32, c_int32, c_void_p,
                                      c_void_p, c_void_p]),
    "gm_train_layernorm_fwd": (ctypes.c_int, [c_void_p, c_void_p, c_void_p, c_int32] + [c_void_p] * 6),
    "gm_train_layernorm_partial_bytes": (c_size_t, [c_int32]),
    "gm_train_layernorm_bwd": (ctypes.c_int, [c_void_p] * 6 + [c_int32] + [c_void_p] * 5),
    "gm_train_layernorm_bwd_static": (ctypes.c_int, [c_void_p] * 6 + [c_int32] + [c_void_p] * 4),
    "gm_train_small_fwd": (ctypes.c_int, [c_void_p, c_int32, c_void_p, c_void_p, c_int32, c_int32, c_int32, c_void_p,
                                          c_void_p]),
    "gm_train_small_partial_bytes": (c_size_t, [c_int32, c_int32, c_int32]),
    "gm_train_small_bwd": (ctypes.c_int, [c_void_p, c_int32, c_void_p, c_int32, c_void_p, c_void_p, c_int32] +
                           [c_void_p] * 5),
    "gm_train_l1_loss": (ctypes.c_int, [c_void_p, c_void_p, c_void_p, c_int32, c_int32, c_void_p, c_void_p, c_void_p,
                                        c_void_p]),
    "gm_adam_step": (ctypes.c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_float, c_float, c_float,
                                    c_float, c_int32, c_void_p]),
    "gm_train_workspace_bytes": (c_size_t, [POINTER(gm_train_desc)]),
    "gm_train_create": (ctypes.c_int, [POINTER(c_void_p), POINTER(gm_train_desc), c_void_p, c_size_t, c_void_p]),
    "gm_train_destroy": (None, [c_void_p]),
    "gm_train_sync_weights": (ctypes.c_int, [c_void_p, c_void_p]),
    "gm_train_num_buckets": (c_int32, [c_void_p]),
    "gm_train_bucket": (ctypes.c_int, [c_void_p, c_int32, POINTER(c_size_t), POINTER(c_size_t)]),
    "gm_train_prediction": (c_void_p, [c_void_p]),
    "gm_train_activation": (c_void_p, [c_void_p, c_int32, c_int32]),
    "gm_train_forward_backward": (ctypes.c_int, [c_void_p, c_void_p, c_void_p, c_int32, c_double, c_int32, c_int32] +
                                  [c_void_p] * 9 + [c_int32, c_void_p, c_void_p, c_void_p]),
    "gm_random_walk_noise": (ctypes.c_int, [c_void_p, c_int32, c_int32, c_int32, POINTER(gm_phys), c_void_p, c_double,
                                            ctypes.c_uint64, c_void_p, c_void_p, c_void_p, c_void_p]),
    "gm_training_target": (ctypes.c_int, [c_void_p, c_int32, c_int32, c_int32, POINTER(gm_phys), c_void_p, c_void_p,
                                          c_void_p]),
    "gm_debug_stamps": (ctypes.c_int, [c_void_p, c_int32]),
    "gm_selftest_gemm": (ctypes.c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int32, c_void_p]),
    "gm_selftest_gemm_ts": (ctypes.c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int32, c_void_p]),
}

_lib = None


def load():
    """Load the shared library (no device needed). Raises if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(make -C gnn_manip_b200/csrc). gnn_manip_b200 has no CPU fallback.")
        lib = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def lib():
    """The library, for compute calls: additionally requires a CUDA device."""
    l = load()
    if not torch.cuda.is_available():
        raise RuntimeError("gnn_manip_b200 needs a CUDA device (sm_100a); there is no CPU path")
    return l


def check(rc, what=""):
    if rc != 0:
        msg = load().gm_last_error().decode(errors="replace")
        raise GmError(f"{what or 'gnn_manip_b200'} failed (status {rc}): {msg}")


def stream_ptr(device=None):
    return c_void_p(torch.cuda.current_stream(device).cuda_stream)


def ptr(t):
    """Device pointer of a tensor (None -> NULL)."""
    return c_void_p(0 if t is None else t.data_ptr())


def launch_count():
    return int(load().gm_launch_count())


def compute_device(t=None):
    """Device the kernels run on for an input living on `t.device` (CPU inputs go to the current GPU)."""
    if t is not None and t.is_cuda:
        return t.device
    if not torch.cuda.is_available():
        raise RuntimeError("gnn_manip_b200 needs a CUDA device (sm_100a); there is no CPU path")
    return torch.device("cuda", torch.cuda.current_device())


def make_phys(conn_r, stats, bounds, cartesian_idx, material_idx, control_idx=None):
    """gm_phys from the reference's metadata dicts (gnn_manip/utils/coffee_dataset.py:18-43)."""
    p = gm_phys()
    mat = material_idx[0] if isinstance(material_idx, (list, tuple)) else int(material_idx)
    p.cartesian_idx[:] = [int(i) for i in cartesian_idx]
    p.material_idx = int(mat)
    p.use_control = 0 if control_idx is None else 1
    p.control_idx[:] = [int(i) for i in (control_idx if control_idx is not None else [0, 0, 0])]
    for name, key in (("velocity_mean", "velocity_mean"), ("velocity_std", "velocity_std"),
                      ("acceleration_mean", "acceleration_mean"), ("acceleration_std", "acceleration_std")):
        v = stats[key] if stats is not None and key in stats else ([0.0, 0.0, 0.0] if "mean" in key else [1.0] * 3)
        getattr(p, name)[:] = [float(x) for x in torch.as_tensor(v, dtype=torch.float32).reshape(-1)[:3]]
    lo = bounds["lower_bounds"] if bounds is not None else [0.0] * 3
    hi = bounds["upper_bounds"] if bounds is not None else [1.0] * 3
    p.lower_bounds[:] = [float(x) for x in torch.as_tensor(lo, dtype=torch.float32).reshape(-1)[:3]]
    p.upper_bounds[:] = [float(x) for x in torch.as_tensor(hi, dtype=torch.float32).reshape(-1)[:3]]
    p.conn_r = float(conn_r)
    return p
