"""The reference's OWN acceptance tests for its in-memory store, run against the drop-in
(adapters/copilot_vectorstore/tests/test_inmemory.py:14-192; SURVEY.md section 8c "Acceptance on the caller side").

The test module is not copied into this repository: oracle/make_ref.py byte-compiles it (and the reference's
``interface`` / ``inmemory`` modules) from /root/reference into the git-ignored oracle/_ref/, which travels to the
GPU box (oracle/ref_loader.py loads it).  Here the compiled test module is imported with ``copilot_vectorstore.inmemory.InMemoryVectorStore``
rebound to ``B200VectorStore`` on cuda:0, and every ``test_*`` method of every ``Test*`` class is run.  With
oracle/_ref on the path the drop-in subclasses the reference's own ``VectorStore`` and returns the reference's own
``SearchResult`` (vectorstore/interface.py), which is what the ``isinstance`` assertions of those tests check."""
import os
import subprocess
import sys

import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu
REFDIR = os.path.join(ROOT, "oracle", "_ref")

_RUN = r"""
import sys, traceback
sys.path.insert(0, {root!r})
import importlib
from oracle import ref_loader
assert ref_loader.load()
import copilot_vectorstore.inmemory as ref_inmemory
import copilot_vectorstore.interface as ref_interface
vs = importlib.import_module("copilot-for-consensus_b200.vectorstore")
assert vs.interface.HAVE_REFERENCE_INTERFACE and issubclass(vs.B200VectorStore, ref_interface.VectorStore)
real = ref_inmemory.InMemoryVectorStore
class DropIn(vs.B200VectorStore):
    def __init__(self):
        super().__init__(device="cuda:0")
ref_inmemory.InMemoryVectorStore = DropIn
assert ref_loader.load(include_tests=True)
import reftests.test_inmemory as t
ref_inmemory.InMemoryVectorStore = real
ran, failed = 0, []
for cname in sorted(n for n in dir(t) if n.startswith("Test")):
    cls = getattr(t, cname)
    for m in sorted(n for n in dir(cls) if n.startswith("test_")):
        ran += 1
        try:
            getattr(cls(), m)()
        except Exception:
            failed.append(cname + "." + m + "\n" + traceback.format_exc())
print("ran", ran, "failed", len(failed))
for f in failed:
    print(f)
sys.exit(1 if failed or ran < 15 else 0)
"""


def test_reference_inmemory_suite_passes_on_the_drop_in():
    if not os.path.exists(os.path.join(REFDIR, "reftests.test_inmemory.bin")):
        pytest.skip("oracle/_ref is not built (python oracle/make_ref.py needs /root/reference)")
    r = subprocess.run([sys.executable, "-c", _RUN.format(root=ROOT)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-6000:] + r.stderr[-3000:]
    assert "failed 0" in r.stdout
