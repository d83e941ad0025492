"""``SearchResult`` / ``VectorStore``: the reference's own classes when ``copilot_vectorstore`` is
installed (so the drop-in *is a* ``copilot_vectorstore.VectorStore``), else structural equivalents
of adapters/copilot_vectorstore/copilot_vectorstore/interface.py:11-126."""
from __future__ import annotations

from abc import ABC, abstractmethod
from dataclasses import dataclass
from typing import Any

try:  # pragma: no cover - depends on the deployment
    from copilot_vectorstore.interface import SearchResult, VectorStore  # type: ignore

    HAVE_REFERENCE_INTERFACE = True
except Exception:
    HAVE_REFERENCE_INTERFACE = False

    @dataclass
    class SearchResult:  # type: ignore[no-redef]
        id: str
        score: float
        vector: list[float]
        metadata: dict[str, Any]

    class VectorStore(ABC):  # type: ignore[no-redef]
        @abstractmethod
        def add_embedding(self, id: str, vector: list[float], metadata: dict[str, Any]) -> None: ...

        @abstractmethod
        def add_embeddings(self, ids: list[str], vectors: list[list[float]], metadatas: list[dict[str, Any]]) -> None: ...

        @abstractmethod
        def query(self, query_vector: list[float], top_k: int = 10) -> list[SearchResult]: ...

        @abstractmethod
        def delete(self, id: str) -> None: ...

        @abstractmethod
        def clear(self) -> None: ...

        @abstractmethod
        def count(self) -> int: ...

        @abstractmethod
        def get(self, id: str) -> SearchResult: ...
