"""Loader for oracle/_ref -- TEST / BENCHMARK INFRASTRUCTURE, not product code.

oracle/make_ref.py byte-compiles a few reference modules from /root/reference into oracle/_ref/*.bin (Python
bytecode; see there for which and why).  ``load()`` registers them in ``sys.modules`` under their real names
(``copilot_vectorstore.inmemory`` ...), with empty stand-in parent packages, and returns True; it returns False
when oracle/_ref has not been built or was compiled by another Python version.  Only tests/, __graft_entry__ and
bench.py's CPU legs may call it."""
from __future__ import annotations

import importlib.machinery
import importlib.util
import os
import sys
import types

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")


def available() -> bool:
    return os.path.exists(os.path.join(REF_DIR, "copilot_vectorstore.inmemory.bin"))


def _ensure_package(name: str) -> None:
    if name not in sys.modules:
        pkg = types.ModuleType(name)
        pkg.__path__ = []          # a package with nothing importable from disk: every submodule comes from load()
        pkg.__doc__ = "stand-in parent package created by oracle/ref_loader.py"
        sys.modules[name] = pkg
        if "." in name:
            parent, _, leaf = name.rpartition(".")
            _ensure_package(parent)
            setattr(sys.modules[parent], leaf, pkg)


def load(include_tests: bool = False) -> bool:
    if not available():
        return False
    from oracle.make_ref import FILES

    try:
        for _src, modname, fname in FILES:
            if modname.startswith("reftests") and not include_tests:
                continue
            if modname in sys.modules:
                continue
            parent, _, leaf = modname.rpartition(".")
            _ensure_package(parent)
            loader = importlib.machinery.SourcelessFileLoader(modname, os.path.join(REF_DIR, fname))
            spec = importlib.util.spec_from_loader(modname, loader)
            mod = importlib.util.module_from_spec(spec)
            sys.modules[modname] = mod
            try:
                loader.exec_module(mod)
            except BaseException:
                del sys.modules[modname]
                raise
            setattr(sys.modules[parent], leaf, mod)
    except Exception:  # noqa: BLE001 - bad magic number (other Python version), truncated file ...
        return False
    return True
