"""Rerank-stage candidate feed: batched id-matrix -> documents fetch.

CPU only.  The checker is (1) a per-query restatement of the reference's fetch semantics
(document_store.py:193-296) written against plain sqlite, and (2) where /root/reference is
mounted, the reference's OWN DocumentStore run on the same database in a subprocess
(tests/_ref_feed_driver.py)."""
import json
import os
import sqlite3
import subprocess
import sys
import threading

import numpy as np
import pytest

from conftest import REFERENCE_SRC, ROOT

from rag_inference_pipeline_b200 import CandidateDocument, CandidateFeed, install


def make_db(tmp_path, n=500, holes=(7, 123), null_like=(11,)):
    """Table with the reference builder's schema (scripts/create_test_docs.py:29-34)."""
    d = tmp_path / "documents"
    d.mkdir(exist_ok=True)
    conn = sqlite3.connect(str(d / "documents.db"))
    conn.execute("CREATE TABLE documents (doc_id INTEGER PRIMARY KEY, title TEXT NOT NULL, "
                 "content TEXT NOT NULL, category TEXT NOT NULL)")
    rows = []
    for i in range(n):
        if i in holes:
            continue
        cat = "" if i in null_like else f"cat{i % 5}"
        rows.append((i, f"Title {i} " + "t" * (i % 40), f"Body {i} " + "lorem ipsum " * (i % 90), cat))
    conn.executemany("INSERT INTO documents VALUES (?, ?, ?, ?)", rows)
    conn.commit()
    conn.close()
    return d


def per_query_restatement(db, ids, truncate):
    """What the reference returns: one SELECT per query, input order, unknown ids dropped."""
    conn = sqlite3.connect(str(db / "documents.db"))
    out = []
    for row in ids:
        found = {}
        for i in set(int(x) for x in row):
            r = conn.execute("SELECT doc_id, title, content, category FROM documents WHERE doc_id = ?",
                             (i,)).fetchone()
            if r:
                found[i] = r
        docs = []
        for i in row:
            if int(i) in found:
                r = found[int(i)]
                t, c = r[1] or "", r[2] or ""
                if truncate:
                    t, c = t[:truncate], c[:truncate]
                docs.append(CandidateDocument(r[0], t, c, r[3] or None))
        out.append(docs)
    conn.close()
    return out


@pytest.mark.parametrize("B,k,truncate", [(1, 10, 512), (32, 10, 64), (257, 100, None), (8, 1, 16)])
def test_feed_matches_per_query_fetch(tmp_path, B, k, truncate):
    db = make_db(tmp_path)
    rng = np.random.default_rng(B * 1000 + k)
    ids = rng.integers(0, 500, size=(B, k)).astype(np.int64)
    ids[0, -1] = -1                      # padding of a short result
    if B > 1:
        ids[1, :] = ids[0, :]            # two queries with the same neighbours
    feed = CandidateFeed(db_path=db / "documents.db", truncate_length=truncate)
    got = feed.fetch(ids)
    want = per_query_restatement(db, ids, truncate)
    assert got == want
    assert feed.stats["selects"] == 1 and feed.stats["ids"] == B * k
    assert feed.stats["distinct_ids"] == np.unique(ids[ids >= 0]).size
    # the DocumentStore-style entry points give the same documents
    assert feed.fetch_documents_batch(ids.tolist(), truncate_length=truncate) == want
    assert feed.fetch_documents(ids[0].tolist()) == per_query_restatement(db, ids[:1], None)[0]
    feed.close_all()


def test_feed_edge_cases(tmp_path):
    db = make_db(tmp_path)
    feed = CandidateFeed(db_path=db / "documents.db", truncate_length=32)
    assert feed.fetch(np.zeros((0, 10), np.int64)) == []
    assert feed.fetch(np.full((3, 4), -1, np.int64)) == [[], [], []]
    assert feed.fetch_documents([]) == [] and feed.fetch_documents_batch([]) == []
    docs = feed.fetch(np.array([[7, 11, 11, 123, 12]], np.int64))[0]     # holes dropped, duplicates kept
    assert [d.doc_id for d in docs] == [11, 11, 12]
    assert docs[0].category is None and docs[2].category == "cat2"      # '' -> None (document_store.py:242)
    assert "category" not in docs[0].to_dict() and docs[2].to_dict()["category"] == "cat2"
    assert len(docs[2].content) <= 32 and docs[2].truncate(4).title == "Titl"
    ragged = feed.fetch([[1, 2, 3], [], [4]])
    assert [[d.doc_id for d in r] for r in ragged] == [[1, 2, 3], [], [4]]
    with pytest.raises(FileNotFoundError):
        CandidateFeed(db_path=tmp_path / "nope.db")
    feed.close_all()


def test_feed_id_only_mode_and_scores(tmp_path):
    db = make_db(tmp_path)
    ids = np.array([[5, 6, -1], [7, 8, 9]], np.int64)
    D = np.array([[0.9, 0.8, -3e38], [0.7, 0.6, 0.5]], np.float32)
    feed = CandidateFeed(db_path=db / "documents.db", payload_mode="id_only")
    got = feed.fetch(ids)                                                # api.py:445-449
    assert [[(d.doc_id, d.title, d.content) for d in r] for r in got] == [
        [(5, "", ""), (6, "", "")], [(7, "", ""), (8, "", ""), (9, "", "")]]
    assert feed.stats["selects"] == 0
    full = CandidateFeed(db_path=db / "documents.db", payload_mode="full", truncate_length=20)
    pairs = full.fetch_with_scores(D, ids)
    assert [(d.doc_id, round(s, 3)) for d, s in pairs[0]] == [(5, 0.9), (6, 0.8)]
    # id 7 is not in the table: its score must not slide onto id 8
    assert [(d.doc_id, round(s, 3)) for d, s in pairs[1]] == [(8, 0.6), (9, 0.5)]
    feed.close_all()
    full.close_all()


def test_feed_from_worker_threads(tmp_path):
    """Thread-local connections over the shared in-memory copy (document_store.py:131-151)."""
    db = make_db(tmp_path)
    feed = CandidateFeed(db_path=db / "documents.db", truncate_length=48)
    rng = np.random.default_rng(3)
    batches = [rng.integers(0, 500, size=(16, 10)).astype(np.int64) for _ in range(8)]
    results = [None] * len(batches)

    def work(i):
        results[i] = feed.fetch(batches[i])

    threads = [threading.Thread(target=work, args=(i,)) for i in range(len(batches))]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    for i, b in enumerate(batches):
        assert results[i] == per_query_restatement(db, b, 48)
    feed.close_all()


def test_install_registers_feed_under_document_store(tmp_path):
    class Key:
        def __init__(self, v):
            self.value = v

    ds = Key("document_store")
    factories = {ds: object, Key("faiss"): object}
    install(factories, replace_document_store=True)
    db = make_db(tmp_path)

    class S:
        documents_dir = str(db)
        documents_payload_mode = "full"
        truncate_length = 8

    feed = factories[ds](S(), {})
    assert isinstance(feed, CandidateFeed)
    assert feed.fetch(np.array([[3]], np.int64))[0][0].content == "Body 3 l"
    feed.close_all()


@pytest.mark.skipif(not os.path.isdir(REFERENCE_SRC), reason="/root/reference not available on this box")
def test_feed_matches_reference_document_store(tmp_path):
    db = make_db(tmp_path, n=800)
    out = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "_ref_feed_driver.py"), str(db)],
                         capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr[-3000:]
    res = json.loads(out.stdout.strip().splitlines()[-1])
    assert res["ok"], res
    assert res["queries"] == 300 and res["selects_feed"] == 1


def test_truncate_length_zero_and_no_aliasing(tmp_path):
    """Reference semantics (document_store.py:297-298): truncate_length=0 empties every field
    (`is not None` test), None leaves the text alone; documents are not shared between queries."""
    db = make_db(tmp_path, 50)
    feed = CandidateFeed(db_path=db / "documents.db", truncate_length=None)
    ids = [[1, 2, 3], [3, 2, 7]]
    empty = feed.fetch_documents_batch(ids, truncate_length=0)
    assert all(d.title == "" and d.content == "" for row in empty for d in row)
    full = feed.fetch_documents_batch(ids, truncate_length=None)
    assert all(len(d.content) > 0 for row in full for d in row)
    assert full[0][2] == full[1][0] and full[0][2] is not full[1][0]
    full[0][2].title = "mutated"
    assert full[1][0].title != "mutated"
    assert feed.fetch_documents([3])[0].title != "mutated"
