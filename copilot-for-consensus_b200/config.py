"""Typed adapter configs.  When the reference's ``copilot_config`` package is importable its
generated dataclasses are used as-is (so ``isinstance`` checks in the reference keep working);
otherwise structurally identical stand-ins are defined here
(adapters/copilot_config/copilot_config/generated/adapters/embedding_backend.py:64-87,
 .../vector_store.py:45-82)."""
from __future__ import annotations

from dataclasses import dataclass
from typing import Any

try:  # pragma: no cover - depends on the deployment
    from copilot_config.generated.adapters.embedding_backend import (  # type: ignore
        AdapterConfig_EmbeddingBackend,
        DriverConfig_EmbeddingBackend_Sentencetransformers,
    )
    from copilot_config.generated.adapters.vector_store import (  # type: ignore
        AdapterConfig_VectorStore,
        DriverConfig_VectorStore_Inmemory,
    )

    HAVE_REFERENCE_CONFIG = True
except Exception:
    HAVE_REFERENCE_CONFIG = False

    @dataclass
    class DriverConfig_EmbeddingBackend_Sentencetransformers:  # type: ignore[no-redef]
        cache_dir: str | None = None
        device: str = "cpu"
        model_name: str = "all-MiniLM-L6-v2"

    @dataclass
    class AdapterConfig_EmbeddingBackend:  # type: ignore[no-redef]
        embedding_backend_type: str
        driver: Any

    @dataclass
    class DriverConfig_VectorStore_Inmemory:  # type: ignore[no-redef]
        pass

    @dataclass
    class AdapterConfig_VectorStore:  # type: ignore[no-redef]
        vector_store_type: str
        driver: Any
