"""Rerank-stage candidate feed: the step right after ``store.search``.

The reference turns the ``[B, k]`` id matrix into ``list[list[int]]`` and then fetches the
documents one QUERY at a time (a thread pool of per-query ``fetch_documents_batch([doc_ids])``
calls, each of which fills a temp table row by row and runs its own SELECT):

    /root/reference/src/pipeline/services/retrieval/api.py:427-473        _fetch_documents
    /root/reference/src/pipeline/components/document_store.py:193-268     fetch_documents
    /root/reference/src/pipeline/components/document_store.py:270-296     fetch_documents_batch

At B = 1024, k = 10 that is 1024 temp-table refills and SELECTs for ~10 k ids of which many are
shared between queries.  ``CandidateFeed`` takes the id matrix as the search operator returns it
(``int64 [B, k]``, ``-1`` padding), fetches every DISTINCT id of the batch once (one temp-table
fill with ``executemany`` over a numpy view, one SELECT), truncates each distinct document once,
and hands back the same structure the reference's executor expects: one list of documents per
query, in id-matrix order, ids that are not in the table silently dropped
(document_store.py:268), ``title``/``content`` never ``None`` (:240-241), empty ``category`` ->
``None`` (:242).

It keeps the reference's DocumentStore interface (``fetch_documents``, ``fetch_documents_batch``,
``clear_cache``, ``close_all``; thread-local connections, shared in-memory copy of the database)
so it can be registered under ``"document_store"`` in the reference's registry.  Host-only code:
sqlite is the storage engine on both sides and stays out of scope.
"""
from __future__ import annotations

import copy
import sqlite3
import threading
from pathlib import Path
from typing import Any, Iterable, Sequence

import numpy as np

__all__ = ["CandidateDocument", "CandidateFeed", "create_candidate_feed"]


CONFIGURED = object()   # sentinel: "use settings.truncate_length" (None means: do not truncate)


class CandidateDocument:
    """Same fields and helpers as the reference's ``Document`` (document_store.py:24-85)."""

    __slots__ = ("doc_id", "title", "content", "category")

    def __init__(self, doc_id: int, title: str, content: str, category: str | None = None) -> None:
        self.doc_id = doc_id
        self.title = title
        self.content = content
        self.category = category

    def to_dict(self) -> dict[str, str | int]:
        out: dict[str, str | int] = {"doc_id": self.doc_id, "title": self.title, "content": self.content}
        if self.category is not None:
            out["category"] = self.category
        return out

    def truncate(self, max_length: int) -> "CandidateDocument":
        return CandidateDocument(self.doc_id, self.title[:max_length] if self.title else "",
                                 self.content[:max_length] if self.content else "", self.category)

    def __eq__(self, other: object) -> bool:
        return (isinstance(other, CandidateDocument) and self.doc_id == other.doc_id
                and self.title == other.title and self.content == other.content
                and self.category == other.category)

    def __repr__(self) -> str:
        return f"CandidateDocument(doc_id={self.doc_id}, title={self.title[:24]!r})"


_SELECT = ("SELECT doc_id, title, content, category FROM documents "
           "WHERE doc_id IN (SELECT doc_id FROM temp_doc_ids)")


class CandidateFeed:
    """Batched document fetch keyed by the search operator's id matrix.

    ``settings`` is the reference's ``PipelineSettings`` (fields read: ``documents_dir``,
    ``documents_payload_mode``, ``truncate_length``); every field can be overridden by keyword so
    the feed also works without the reference's config package.
    """

    def __init__(self, settings: Any = None, config: dict[str, Any] | None = None, *,
                 db_path: str | Path | None = None, payload_mode: str | None = None,
                 truncate_length: int | None = None, in_memory: bool | None = None) -> None:
        config = config or {}
        self.settings = settings
        if db_path is None:
            db_path = config.get("db_path")
        if db_path is None:
            db_path = Path(getattr(settings, "documents_dir", "documents/")) / "documents.db"
        self.db_path = Path(db_path)
        self.payload_mode = payload_mode or config.get("payload_mode") or getattr(
            settings, "documents_payload_mode", "full")
        self.truncate_length = truncate_length if truncate_length is not None else config.get(
            "truncate_length", getattr(settings, "truncate_length", None))
        if not self.db_path.exists():   # document_store.py:111-113
            raise FileNotFoundError(f"Document database not found at {self.db_path}")
        self._local = threading.local()
        self._lock = threading.Lock()
        self._memory_uri = f"file:candidate_feed_{id(self):x}?mode=memory&cache=shared"
        self._memory_conn: sqlite3.Connection | None = None
        self._connections: list[sqlite3.Connection] = []
        if in_memory is None:
            in_memory = bool(config.get("in_memory", self.payload_mode != "id_only"))
        if in_memory:
            self._clone_into_memory()
        self.stats = {"batches": 0, "ids": 0, "distinct_ids": 0, "selects": 0}

    # ------------------------------------------------------------------ connections
    def _clone_into_memory(self) -> None:
        """document_store.py:166-191: keep the whole table in RAM, shared by all threads."""
        with self._lock:
            if self._memory_conn is not None:
                return
            disk = sqlite3.connect(str(self.db_path))
            mem = sqlite3.connect(self._memory_uri, uri=True, check_same_thread=False)
            try:
                disk.backup(mem)
                self._memory_conn = mem
            except sqlite3.Error:
                mem.close()
            finally:
                disk.close()

    def _connection(self) -> sqlite3.Connection:
        conn = getattr(self._local, "conn", None)
        if conn is None:
            if self._memory_conn is not None:
                conn = sqlite3.connect(self._memory_uri, uri=True, check_same_thread=False)
            else:
                conn = sqlite3.connect(str(self.db_path), check_same_thread=False)
            conn.execute("CREATE TEMP TABLE IF NOT EXISTS temp_doc_ids(doc_id INTEGER PRIMARY KEY)")
            self._local.conn = conn
            with self._lock:
                self._connections.append(conn)
        return conn

    # ------------------------------------------------------------------ the batched fetch
    def _fetch_distinct(self, distinct: Iterable[int]) -> dict[int, CandidateDocument]:
        conn = self._connection()
        cur = conn.cursor()
        try:
            cur.execute("DELETE FROM temp_doc_ids")
            cur.executemany("INSERT OR IGNORE INTO temp_doc_ids(doc_id) VALUES (?)",
                            ((int(i),) for i in distinct))
            cur.execute(_SELECT)
            rows = cur.fetchall()
        except sqlite3.Error as e:   # document_store.py:257-260
            raise RuntimeError(f"Failed to fetch documents: {e}") from e
        self.stats["selects"] += 1
        return {r[0]: CandidateDocument(r[0], r[1] or "", r[2] or "", r[3] or None) for r in rows}

    def fetch(self, ids: np.ndarray | Sequence[Sequence[int]],
              truncate_length: int | None | object = CONFIGURED) -> list[list[CandidateDocument]]:
        """ids ``[B, k]`` (or ragged lists) -> one document list per query, id-matrix order.

        ``truncate_length``: ``CONFIGURED`` (default) uses ``settings.truncate_length``; ``None``
        leaves the text untouched; an integer truncates every field to that many characters --
        ``0`` yields empty strings, exactly like the reference (``document_store.py:297-298``
        tests ``is not None``).  Documents are copied per query: mutating one result list never
        shows up in another.
        """
        if isinstance(ids, np.ndarray):
            mat = np.ascontiguousarray(ids, dtype=np.int64)
            if mat.ndim == 1:
                mat = mat[None, :]
            per_query: list[Any] = list(mat)
            flat = mat.ravel()
        else:
            per_query = [np.asarray(r, dtype=np.int64).ravel() for r in ids]
            flat = np.concatenate(per_query) if per_query else np.zeros(0, np.int64)
        self.stats["batches"] += 1
        self.stats["ids"] += int(flat.size)
        if self.payload_mode == "id_only":   # api.py:445-449: ids travel, text does not
            return [[CandidateDocument(int(i), "", "") for i in row if i >= 0] for row in per_query]
        distinct = np.unique(flat[flat >= 0])
        self.stats["distinct_ids"] += int(distinct.size)
        docs = self._fetch_distinct(distinct.tolist()) if distinct.size else {}
        if truncate_length is CONFIGURED:
            truncate_length = self.truncate_length
        if truncate_length is not None:
            docs = {i: d.truncate(int(truncate_length)) for i, d in docs.items()}
        out = []
        for row in per_query:
            out.append([copy.copy(docs[i]) for i in row.tolist() if i in docs])
        return out

    def fetch_with_scores(self, distances: np.ndarray, ids: np.ndarray,
                          truncate_length: int | None | object = CONFIGURED,
                          ) -> list[list[tuple[CandidateDocument, float]]]:
        """(D, I) of ``store.search`` -> per query ``[(document, score), ...]`` for the reranker
        (the pairing api.py:483-492 does with ``zip(docs, scores)``, except that a score stays with
        ITS id when an id is missing from the table)."""
        D = np.asarray(distances)
        mat = np.ascontiguousarray(ids, dtype=np.int64)
        docs = self.fetch(mat, truncate_length)
        out = []
        for q, row_docs in enumerate(docs):
            score_of = {}
            for j, i in enumerate(mat[q].tolist()):
                score_of.setdefault(i, float(D[q, j]))
            out.append([(d, score_of[d.doc_id]) for d in row_docs])
        return out

    # ------------------------------------------------------------------ reference-compatible surface
    def fetch_documents(self, doc_ids: list[int]) -> list[CandidateDocument]:
        """document_store.py:193: documents of one id list, input order, untruncated."""
        if not doc_ids:
            return []
        return self.fetch([doc_ids], truncate_length=None)[0]

    def fetch_documents_batch(self, doc_ids_batch: list[list[int]], truncate_length: int | None = None,
                              ) -> list[list[CandidateDocument]]:
        """document_store.py:270: same result, one SELECT for the whole batch."""
        if not doc_ids_batch:
            return []
        return self.fetch(doc_ids_batch, truncate_length=truncate_length)

    def clear_cache(self) -> None:
        """Nothing cached per id: the in-memory database IS the cache."""

    def close_all(self) -> None:
        with self._lock:
            for c in self._connections:
                try:
                    c.close()
                except sqlite3.Error:
                    pass
            self._connections.clear()
            if self._memory_conn is not None:
                self._memory_conn.close()
                self._memory_conn = None
        self._local = threading.local()

    def __repr__(self) -> str:
        return f"CandidateFeed(db_path={self.db_path}, mode={self.payload_mode})"


def create_candidate_feed(settings: Any, config: dict[str, Any] | None = None) -> CandidateFeed:
    """Factory with the reference's signature (component_factory.py:30-31)."""
    return CandidateFeed(settings, config or {})
