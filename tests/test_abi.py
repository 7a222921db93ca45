"""CPU suite: the C-ABI boundary.  The shared library loads without a GPU and exports exactly the symbols that
include/gnn_manip_b200.h declares; the ctypes table in gnn_manip_b200/_lib.py covers every one of them; the product
package never touches oracle/ and has no CPU fallback (compute entry points raise without a CUDA device)."""
import ctypes
import os
import re
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "gnn_manip_b200.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(gm_[a-z0-9_]+)\s*\(", src)))


@pytest.fixture(scope="module")
def lib():
    subprocess.check_call(["make", "-s", "-j", "8", "-C", os.path.join(ROOT, "gnn_manip_b200", "csrc")])
    from gnn_manip_b200 import _lib
    return _lib


def test_library_exports_every_declared_symbol(lib):
    L = lib.load()
    names = declared_symbols()
    assert len(names) >= 30
    for n in names:
        assert hasattr(L, n), f"{n} is declared in include/gnn_manip_b200.h but not exported"
    assert L.gm_version() >= 100


def test_ctypes_table_matches_header(lib):
    assert sorted(lib.SIGNATURES) == declared_symbols()


def test_exported_symbols_are_plain_c(lib):
    out = subprocess.run(["nm", "-D", "--defined-only", lib.LIB_PATH], capture_output=True, text=True).stdout
    exported = {l.split()[-1] for l in out.splitlines() if " T " in l}
    assert set(declared_symbols()) <= exported


def test_pure_queries_work_without_a_gpu(lib):
    L = lib.load()
    assert L.gm_node_dim(6, 1) == 25 and L.gm_node_dim(6, 0) == 22
    assert L.gm_model_num_params(25, 4, 3, 128, 2, 10) == 1591299
    assert L.gm_model_blob_bytes(25, 4, 3, 128, 2, 10) > 1591299 * 4
    assert L.gm_radius_graph_workspace_bytes(2000, 1, 20) > 0
    assert L.gm_radius_graph_workspace_bytes(0, 1, 20) == 0


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU behaviour")
def test_no_cpu_fallback(lib):
    import gnn_manip_b200 as gm
    with pytest.raises(RuntimeError, match="CUDA device"):
        gm.get_connectivity(torch.rand(10, 3), 0.015)
    m = gm.EncProcDecGNN(25, 4, 3, 128, 2, 2)
    with pytest.raises(RuntimeError, match="CUDA device"):
        m(torch.zeros(4, 25), torch.zeros(2, 4), torch.zeros(2, 2, dtype=torch.long))


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "gnn_manip_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", text, flags=re.M), f"{f} imports oracle"
                assert "restate" not in text, f"{f} mentions the oracle restatement"
    code = "import sys; import gnn_manip_b200; sys.exit(1 if any(m.split('.')[0]=='oracle' for m in sys.modules) else 0)"
    assert subprocess.run([sys.executable, "-c", code], cwd=ROOT).returncode == 0


def test_state_dict_layout_of_the_drop_in_model():
    import gnn_manip_b200 as gm
    sd = gm.EncProcDecGNN(25, 4, 3, 128, 2, 10).state_dict()
    assert len(sd) == 342
    assert sd["processor.3.phi_edge.0.weight"].shape == (128, 384)
    assert sd["processor.3.edge_model.phi_edge.0.weight"].data_ptr() == sd["processor.3.phi_edge.0.weight"].data_ptr()
    assert sd["processor.9.node_model.phi_node.0.weight"].shape == (128, 256)
    assert sd["encoder.phi_node.0.weight"].shape == (128, 25) and sd["decoder.4.weight"].shape == (3, 128)
    unique = {v.data_ptr(): v.numel() for v in sd.values()}
    assert len(unique) == 182 and sum(unique.values()) == 1591299
