"""``create_embedding_provider`` with the reference's signature and error contract
(adapters/copilot_embedding/copilot_embedding/factory.py:26-76).

Routing happens at the factory level only: the ``sentencetransformers``
driver with ``device`` matching ``cuda(:N)?`` is served by the B200 provider; every other
combination (including ``device="cpu"``) goes to the reference's own factory when
``copilot_embedding`` is installed, and is an unknown driver otherwise."""
from __future__ import annotations

import re
from typing import Any

from .base import EmbeddingProvider

_CUDA = re.compile(r"^cuda(:\d+)?$")
_REFERENCE_BACKENDS = ("mock", "sentencetransformers", "openai", "azure_openai", "huggingface")


def create_embedding_provider(config: Any) -> EmbeddingProvider:
    if config is None:
        raise ValueError("embedding_backend config is required")
    backend = str(config.embedding_backend_type).lower()
    driver_config = config.driver
    if backend == "sentencetransformers":
        if type(driver_config).__name__ != "DriverConfig_EmbeddingBackend_Sentencetransformers":
            raise TypeError("driver_config must be DriverConfig_EmbeddingBackend_Sentencetransformers")
        if _CUDA.match(str(driver_config.device)):
            from .provider import B200SentenceTransformerProvider

            return B200SentenceTransformerProvider.from_config(driver_config)
    if backend in _REFERENCE_BACKENDS:
        try:
            from copilot_embedding.factory import create_embedding_provider as reference_factory  # type: ignore
        except Exception as exc:
            raise ValueError(
                f"Unknown embedding backend driver: {backend} on this deployment. Supported backends: "
                f"sentencetransformers with device cuda[:N] (the reference's copilot_embedding package is not "
                f"installed: {exc})"
            ) from exc
        return reference_factory(config)
    raise ValueError(
        f"Unknown embedding backend driver: {backend}. "
        f"Supported backends: mock, sentencetransformers, openai, azure_openai, huggingface"
    )
