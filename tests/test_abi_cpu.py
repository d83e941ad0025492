"""CPU-side checks of the C-ABI boundary: the library loads, exports exactly the symbols
include/b200_retrieval.h declares, and fails loudly (no fallback) when there is no GPU."""
import ctypes
import os
import re

import pytest

from conftest import ROOT, pkg


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "b200_retrieval.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    nat = pkg("_native")
    if not os.path.exists(nat.LIB_PATH):
        import __graft_entry__ as ge

        ge.build()
    L = ctypes.CDLL(nat.LIB_PATH)
    declared = _declared_symbols()
    assert len(declared) >= 20
    for name in declared:
        assert hasattr(L, name), f"{name} declared in the header but not exported"
    assert sorted(nat.SYMBOLS) == declared, "ctypes table and header disagree"


def test_no_gpu_means_loud_failure():
    import torch

    if torch.cuda.is_available():
        pytest.skip("this check is for the GPU-less build container")
    nat = pkg("_native")
    L = nat.lib()
    assert L.b200_device_count() == 0
    h = ctypes.c_void_p()
    rc = L.b200_index_create(384, 1024, 0, 0, ctypes.byref(h))
    assert rc != 0 and not h.value
    with pytest.raises(RuntimeError, match="no CPU fallback|CUDA"):
        nat.check(rc)
    vs = pkg("vectorstore")
    st = vs.B200VectorStore(device="cuda:0")
    with pytest.raises(RuntimeError):
        st.add_embedding("a", [1.0, 0.0], {})


def test_status_to_exception_mapping():
    nat = pkg("_native")
    nat.lib()
    for rc, exc in [(-1, ValueError), (-2, ValueError), (-3, KeyError), (-4, RuntimeError), (-5, RuntimeError), (-6, MemoryError)]:
        with pytest.raises(exc):
            nat.check(rc)
    nat.check(0)
    assert nat.device_index("cuda:3") == 3 and nat.device_index(2) == 2
    with pytest.raises(ValueError):
        nat.device_index("cpu")


def test_alias_module_imports_the_package():
    import copilot_for_consensus_b200 as alias

    assert alias is pkg()
    assert hasattr(alias, "create_vector_store")
