"""``create_vector_store`` with the reference's signature and error contract
(adapters/copilot_vectorstore/copilot_vectorstore/factory.py:57-78 via
adapters/copilot_config/copilot_config/adapter_factory.py:26-68).

Routing happens at the factory level only: the ``inmemory`` driver slot is
served by :class:`B200VectorStore`; every other driver type is handed to the reference's own factory
when ``copilot_vectorstore`` is installed, and is an unknown driver otherwise."""
from __future__ import annotations

from typing import Any

from .interface import VectorStore
from .store import B200VectorStore

_REFERENCE_DRIVERS = ("azure_ai_search", "faiss", "inmemory", "qdrant")

#: constructor arguments for stores built through the factory; the inmemory driver schema has no
#: properties (docs/schemas/configs/adapters/drivers/vector_store/vectorstore_inmemory.json), and the
#: adapters are test-pinned not to read the environment, so deployment settings live here.
STORE_DEFAULTS: dict[str, Any] = {"device": "cuda:0"}


#: opt-in: also serve the ``faiss`` driver slot when it asks for a flat index without persistence
#: (faiss_store.py:99-101: ``IndexFlatL2``) with the B200 store in ``metric="l2"`` mode
SERVE_FAISS_FLAT = {"enabled": False}


def configure(serve_faiss_flat: bool | None = None, **kwargs: Any) -> None:
    """Set constructor defaults (``device``, ``upsert``, ``initial_capacity``) for factory-built stores."""
    if serve_faiss_flat is not None:
        SERVE_FAISS_FLAT["enabled"] = bool(serve_faiss_flat)
    STORE_DEFAULTS.update(kwargs)


def _validate_against_schema(driver_type: str, driver: Any) -> None:
    """adapter_factory.py:61-67: the reference validates the driver config against its JSON schema before
    building the driver.  Done with the reference's own validator when ``copilot_config`` is installed (the
    schema files live in that package's tree); without it there is nothing to validate against."""
    try:
        from copilot_config.schema_validation import validate_driver_config_against_schema  # type: ignore
    except Exception:
        return
    validate_driver_config_against_schema(adapter="vector_store", driver=driver_type, config=driver)


def create_vector_store(config: Any) -> VectorStore:
    if config is None:
        raise ValueError("vector_store config is required")
    driver_type = str(config.vector_store_type).lower()
    if driver_type == "inmemory":
        driver = config.driver
        _validate_against_schema(driver_type, driver)
        if type(driver).__name__ != "DriverConfig_VectorStore_Inmemory":
            raise TypeError(f"Expected inmemory config, got {type(driver).__name__}")
        return B200VectorStore.from_config(driver, **STORE_DEFAULTS)
    if driver_type == "faiss" and SERVE_FAISS_FLAT["enabled"]:
        driver = config.driver
        if (type(driver).__name__ == "DriverConfig_VectorStore_Faiss" and getattr(driver, "index_type", "flat") == "flat"
                and not getattr(driver, "persist_path", None)):
            _validate_against_schema(driver_type, driver)
            return B200VectorStore(metric="l2", dimension=int(driver.dimension), **STORE_DEFAULTS)
    if driver_type in _REFERENCE_DRIVERS:
        try:
            from copilot_vectorstore.factory import create_vector_store as reference_factory  # type: ignore
        except Exception as exc:
            raise ValueError(
                f"Unknown vector_store driver: {driver_type}. Supported drivers: inmemory "
                f"(the reference's copilot_vectorstore package is not installed: {exc})"
            ) from exc
        return reference_factory(config)
    supported = ", ".join(sorted(_REFERENCE_DRIVERS))
    raise ValueError(f"Unknown vector_store driver: {driver_type}. Supported drivers: {supported}")
