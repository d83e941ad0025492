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


def test_options_replace_environment_reads():
    """Kernel-variant switches go through b200_set_option; nothing under csrc/ reads the environment
    (the reference pins its adapters to that: adapters/copilot_vectorstore/tests/test_factory.py:75-107)."""
    nat = pkg("_native")
    assert nat.get_option("fuse_ffn") == 1 and nat.get_option("pdl") == 0 and nat.get_option("search_certify") == 1
    with nat.options(fuse_ffn=0, sim_bound=1):
        assert nat.get_option("fuse_ffn") == 0 and nat.get_option("sim_bound") == 1
    assert nat.get_option("fuse_ffn") == 1 and nat.get_option("sim_bound") == 3
    with pytest.raises(ValueError, match="unknown option"):
        nat.set_option("no_such_switch", 1)
    with pytest.raises(ValueError, match="takes"):
        nat.set_option("pdl", 7)
    csrc = os.path.join(ROOT, "copilot-for-consensus_b200", "csrc")
    for fn in os.listdir(csrc):
        if fn.endswith((".cu", ".cuh", ".h")):
            assert "getenv" not in open(os.path.join(csrc, fn)).read(), f"{fn} reads the environment"


def test_alias_module_imports_the_package():
    import copilot_for_consensus_b200 as alias

    assert alias is pkg()
    assert hasattr(alias, "create_vector_store")
