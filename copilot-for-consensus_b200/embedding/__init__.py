"""Drop-in surface of ``copilot_embedding`` (adapters/copilot_embedding/copilot_embedding/__init__.py:34-44)."""
from .base import EmbeddingProvider
from .factory import create_embedding_provider

__all__ = ["EmbeddingProvider", "create_embedding_provider", "B200SentenceTransformerProvider"]


def __getattr__(name):
    if name == "B200SentenceTransformerProvider":
        from .provider import B200SentenceTransformerProvider

        return B200SentenceTransformerProvider
    raise AttributeError(name)
