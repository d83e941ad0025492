"""Import alias: ``import copilot_for_consensus_b200`` -> the package in ``copilot-for-consensus_b200/``
(whose directory name, fixed by the project layout, is not a valid identifier)."""
import importlib
import sys

sys.modules[__name__] = importlib.import_module("copilot-for-consensus_b200")
