"""Drop-in surface of ``copilot_vectorstore`` (adapters/copilot_vectorstore/copilot_vectorstore/__init__.py:11-18)."""
from .factory import configure, create_vector_store
from .interface import SearchResult, VectorStore
from .index import B200Index
from .store import B200VectorStore

__all__ = ["VectorStore", "SearchResult", "create_vector_store", "B200VectorStore", "B200Index", "configure"]
