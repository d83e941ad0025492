"""Build oracle/_ref: the reference's own in-memory vector store, COMPILED (Python bytecode) from the sources
where they lie under /root/reference -- test / benchmark infrastructure, never product code.

    python oracle/make_ref.py            (also run by __graft_entry__.build() when /root/reference exists)

Why: /root/reference does not exist on the GPU box, and the CPU arm of bench.py should time the REAL
``InMemoryVectorStore.query`` (adapters/copilot_vectorstore/copilot_vectorstore/inmemory.py:78-128), not a port
of it.  No reference source is copied anywhere: each listed file is byte-compiled in place with ``py_compile``
and only the resulting bytecode lands in oracle/_ref/ (git-ignored; it travels to the GPU box like the built
``.so`` files).  The files carry the extension ``.bin`` -- the snapshot that goes to the GPU box drops ``*.pyc`` --
and are loaded by oracle/ref_loader.py, which also supplies the (empty) parent packages: the reference's own
``copilot_vectorstore/__init__.py`` imports every driver and their third-party clients and is not used.

What is compiled:
  copilot_vectorstore/interface.py, inmemory.py                 the store and its result type
  copilot_config/generated/adapters/vector_store.py             the driver-config dataclass inmemory.py imports
  copilot_vectorstore/tests/test_inmemory.py                    the reference's own acceptance tests for that store
                                                                (run against the drop-in by tests/test_reference_acceptance_gpu.py)

Only tests/, __graft_entry__ and bench.py's CPU legs may import anything from oracle/_ref.
"""
from __future__ import annotations

import os
import py_compile
import sys

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")

#: (source under /root/reference, module name it is loaded as, file under oracle/_ref)
FILES = [
    ("adapters/copilot_config/copilot_config/generated/adapters/vector_store.py",
     "copilot_config.generated.adapters.vector_store", "copilot_config.generated.adapters.vector_store.bin"),
    ("adapters/copilot_vectorstore/copilot_vectorstore/interface.py", "copilot_vectorstore.interface", "copilot_vectorstore.interface.bin"),
    ("adapters/copilot_vectorstore/copilot_vectorstore/inmemory.py", "copilot_vectorstore.inmemory", "copilot_vectorstore.inmemory.bin"),
    ("adapters/copilot_vectorstore/tests/test_inmemory.py", "reftests.test_inmemory", "reftests.test_inmemory.bin"),
]


def main() -> int:
    if not os.path.isdir(REF):
        print(f"oracle/make_ref.py: {REF} is not here; nothing to build (the prebuilt oracle/_ref is used when present)")
        return 0
    os.makedirs(OUT, exist_ok=True)
    for src, _mod, dst in FILES:
        py_compile.compile(os.path.join(REF, src), cfile=os.path.join(OUT, dst), dfile=src, doraise=True,
                           invalidation_mode=py_compile.PycInvalidationMode.UNCHECKED_HASH)
    with open(os.path.join(OUT, "BUILT_FROM"), "w") as f:
        f.write(f"python {sys.version.split()[0]}; compiled from {REF}:\n" + "".join(f"  {s}\n" for s, _m, _d in FILES))
    print("oracle/_ref built:", ", ".join(d for _, _m, d in FILES))
    return 0


if __name__ == "__main__":
    sys.exit(main())
