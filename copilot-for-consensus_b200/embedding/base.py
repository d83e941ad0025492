"""``EmbeddingProvider``: the reference's own ABC when ``copilot_embedding`` is installed (so the
drop-in *is a* ``copilot_embedding.EmbeddingProvider``), else a structural equivalent of
adapters/copilot_embedding/copilot_embedding/base.py:12-25."""
from __future__ import annotations

from abc import ABC, abstractmethod

try:  # pragma: no cover - depends on the deployment
    from copilot_embedding.base import EmbeddingProvider  # type: ignore

    HAVE_REFERENCE_BASE = True
except Exception:
    HAVE_REFERENCE_BASE = False

    class EmbeddingProvider(ABC):  # type: ignore[no-redef]
        @abstractmethod
        def embed(self, text: str) -> list[float]: ...
