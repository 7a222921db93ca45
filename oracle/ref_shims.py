"""Import shims so the UNMODIFIED reference (/root/reference) runs in the build container.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  Used by oracle/make_golden.py and by
the ``not gpu`` tests that re-validate the restatement when /root/reference is present.
Never imported on the GPU box (the reference does not travel).

Shim set:
  1. torch_scatter.scatter_add(src, index, dim, out)  -> out.index_add_(dim, index, src)
     (called from TorchScript at deps/torch-graphnet/torch_graphnet/utils/gn_utils.py:60-64),
     plus a dummy scatter_mean (graph_networks.py:11).
  2. torch_geometric.nn.MetaLayer, torch_geometric.data.{Data,Dataset,DataLoader} dummies
     (graph_networks.py:12, coffee_dataset.py:9).
  3. gnn_manip.models.action_networks stub (imported by epd_gnn.py:8, absent from the reference).
  4. empty matplotlib / cma / geomloss (gnn_manip/utils/__init__.py:5-7 imports them eagerly).
"""
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("GNN_MANIP_REFERENCE", "/root/reference")


def reference_available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "gnn_manip"))


def _module(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    sys.modules[name] = m
    return m


_installed = False


def install():
    """Install the shims and put the reference on sys.path. Idempotent."""
    global _installed
    if _installed:
        return
    if not reference_available():
        raise RuntimeError(f"reference checkout not found at {REFERENCE_ROOT}")
    import torch

    # (1) torch_scatter
    if "torch_scatter" not in sys.modules:
        @torch.jit.script
        def scatter_add(src: torch.Tensor, index: torch.Tensor, dim: int, out: torch.Tensor):
            return out.index_add_(dim, index, src)

        def scatter_mean(*a, **k):  # never called on the hot path
            raise NotImplementedError

        _module("torch_scatter", scatter_add=scatter_add, scatter_mean=scatter_mean)

    # (2) torch_geometric
    if "torch_geometric" not in sys.modules:
        class _Dummy:
            def __init__(self, *a, **k):
                pass

        class Data(_Dummy):
            def __init__(self, **k):
                self.__dict__.update(k)

        tg = _module("torch_geometric")
        tg.nn = _module("torch_geometric.nn", MetaLayer=_Dummy)
        tg.data = _module("torch_geometric.data", Data=Data, Dataset=_Dummy, DataLoader=_Dummy)

    # (4) plotting / optimiser / loss leaves
    if "matplotlib" not in sys.modules:
        class _Anything:
            """Absorbs any attribute access / call made while importing plot_utils.py."""
            def __getattr__(self, name):
                return _Anything()

            def __call__(self, *a, **k):
                return _Anything()

            def __setitem__(self, k, v):
                pass

        class _PermissiveModule(types.ModuleType):
            def __getattr__(self, name):
                if name.startswith("__"):
                    raise AttributeError(name)
                return _Anything()

        for name in ("matplotlib", "matplotlib.pyplot", "matplotlib.cm", "matplotlib.colors",
                     "mpl_toolkits", "mpl_toolkits.mplot3d"):
            sys.modules[name] = _PermissiveModule(name)
        sys.modules["matplotlib"].rcParams = {}
        sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
    if "cma" not in sys.modules:
        es = _module("cma.evolution_strategy", CMAOptions=dict)
        _module("cma", evolution_strategy=es)
    if "geomloss" not in sys.modules:
        class SamplesLoss:
            def __init__(self, *a, **k):
                pass

            def __call__(self, a, b):
                import torch as _t
                return _t.zeros(())

        _module("geomloss", SamplesLoss=SamplesLoss)

    for p in (REFERENCE_ROOT, os.path.join(REFERENCE_ROOT, "deps", "torch-graphnet")):
        if p not in sys.path:
            sys.path.insert(0, p)

    # (3) missing reference module
    if "gnn_manip.models.action_networks" not in sys.modules:
        import importlib
        pkg = importlib.import_module("gnn_manip")  # noqa: F841  (plain package, empty __init__)
        _module("gnn_manip.models.action_networks",
                ActionInteractionNetwork=object, ActionGraphIndependent=object)
    _installed = True


def load():
    """Return a namespace with the reference symbols on the hot path."""
    install()
    import contextlib
    import io
    with contextlib.redirect_stdout(io.StringIO()):
        from gnn_manip.models.epd_gnn import EncProcDecGNN
        from gnn_manip.utils import utils as ref_utils
        from gnn_manip.utils import collate_utils as ref_collate
        from gnn_manip.utils import rollout_utils as ref_rollout
        from gnn_manip.utils import traj_utils as ref_traj
        import torch_graphnet
    return types.SimpleNamespace(
        EncProcDecGNN=EncProcDecGNN, utils=ref_utils, collate=ref_collate,
        rollout=ref_rollout, traj=ref_traj, torch_graphnet=torch_graphnet)
