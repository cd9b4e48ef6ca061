"""Registry / factory glue: how the store drops into the reference's component system.

The reference builds components through ``COMPONENT_FACTORIES[type](settings, config)``
(``/root/reference/src/pipeline/component_factory.py:17,26-27,63-104``) and the retrieval service
fetches the store with ``registry.get("faiss_store")``
(``/root/reference/src/pipeline/services/retrieval/api.py:218,383,537``).  Two ways to drop in:

1. ``install(COMPONENT_FACTORIES)`` replaces the ``ComponentType.FAISS`` / ``"faiss_store"``
   factories, so the stock profiles (``configs/retrieval_faiss_only.yaml``,
   ``configs/retrieval_embedder_faiss_ids.yaml``) run unchanged;
2. a profile may name the component type ``"flat_ip_b200"`` while keeping the component *name*
   ``faiss_store``.
"""

from __future__ import annotations

from typing import Any

from .store import B200FlatIPStore

COMPONENT_TYPE = "flat_ip_b200"


def create_b200_flat_ip_store(settings, config: dict[str, Any] | None = None) -> B200FlatIPStore:
    """Same signature as the reference's ``create_faiss_store`` (component_factory.py:26-27)."""
    return B200FlatIPStore(settings, config or {})


def install(component_factories: dict, *, replace_faiss: bool = True,
            replace_document_store: bool = False) -> dict:
    """Register the B200 store in a ``COMPONENT_FACTORIES``-style mapping (mutates and returns it).

    ``replace_document_store``: also put the batched ``CandidateFeed`` (candidate_feed.py) behind
    the reference's ``ComponentType.DOCUMENT_STORE`` / ``"document_store"`` key
    (component_factory.py:30-31,68)."""
    component_factories[COMPONENT_TYPE] = create_b200_flat_ip_store
    if replace_document_store:
        from .candidate_feed import create_candidate_feed
        for key in list(component_factories.keys()):
            if key == "document_store" or getattr(key, "value", None) == "document_store":
                component_factories[key] = create_candidate_feed
        component_factories.setdefault("document_store", create_candidate_feed)
    if replace_faiss:
        for key in list(component_factories.keys()):
            if key == "faiss_store" or getattr(key, "value", None) == "faiss":
                component_factories[key] = create_b200_flat_ip_store
        component_factories.setdefault("faiss_store", create_b200_flat_ip_store)
    return component_factories
