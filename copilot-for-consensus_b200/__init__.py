"""copilot-for-consensus_b200 -- the B200-native retrieval hot path of Copilot-for-Consensus.

Chunk embedding (BERT-class encoder) and brute-force cosine top-k, as hand-written sm_100a CUDA
behind a C ABI (``include/b200_retrieval.h``), exposed through the reference's own adapter
surfaces: ``copilot_embedding.create_embedding_provider`` / ``EmbeddingProvider`` and
``copilot_vectorstore.create_vector_store`` / ``VectorStore`` / ``SearchResult``.

The directory name is not a Python identifier; import it with
``importlib.import_module("copilot-for-consensus_b200")`` or through the top-level alias module
``copilot_for_consensus_b200``.
"""
from .embedding import EmbeddingProvider, create_embedding_provider
from .vectorstore import B200VectorStore, SearchResult, VectorStore, create_vector_store

__all__ = ["B200VectorStore", "SearchResult", "VectorStore", "create_vector_store",
           "EmbeddingProvider", "create_embedding_provider"]
__version__ = "0.1.0"
