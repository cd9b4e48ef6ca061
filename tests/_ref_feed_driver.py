"""Runs the REFERENCE's own DocumentStore (imported read-only from /root/reference/src, third-party
imports stubbed as in _ref_integration_driver.py) and CandidateFeed on the same sqlite database and
the same id matrix, and compares every document field.  Also times both (reported, not asserted).
Own process: the reference reads its settings from the environment at import time."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
REF_SRC = "/root/reference/src"


def main(docs_dir):
    import numpy as np

    from _ref_integration_driver import install_stubs
    from rag_inference_pipeline_b200 import CandidateFeed

    os.environ.update({"NODE_NUMBER": "1", "DOCUMENTS_DIR": docs_dir, "DOCUMENTS_PAYLOAD_MODE": "full",
                       "DISABLE_CACHE_FOR_PROFILING": "true"})
    install_stubs()
    sys.path.insert(0, REF_SRC)
    from pipeline.components.document_store import DocumentStore
    from pipeline.config import get_settings

    settings = get_settings()
    ref = DocumentStore(settings)
    feed = CandidateFeed(settings)
    rng = np.random.default_rng(11)
    B, k = 300, 10
    ids = rng.integers(0, 820, size=(B, k)).astype(np.int64)     # some ids beyond the table
    ids[5, 7:] = -1
    ids[6] = ids[4]
    warm = rng.integers(0, 820, size=(4, k)).astype(np.int64)    # connections, lazy imports
    ref.fetch_documents_batch(warm.tolist(), truncate_length=settings.truncate_length)
    feed.fetch(warm)
    feed.stats["selects"] = 0
    t0 = time.perf_counter()
    want = ref.fetch_documents_batch(ids.tolist(), truncate_length=settings.truncate_length)
    t1 = time.perf_counter()
    got = feed.fetch(ids)
    t2 = time.perf_counter()
    selects = feed.stats["selects"]
    same = len(got) == len(want)
    for g, w in zip(got, want):
        same = same and len(g) == len(w) and all(
            a.doc_id == b.doc_id and a.title == b.title and a.content == b.content
            and a.category == b.category and a.to_dict() == b.to_dict() for a, b in zip(g, w))
    one = ref.fetch_documents(ids[0].tolist())
    mine = feed.fetch_documents(ids[0].tolist())
    same = same and [d.to_dict() for d in one] == [d.to_dict() for d in mine]
    print(json.dumps({"ok": bool(same), "queries": B, "selects_feed": selects,
                      "truncate_length": settings.truncate_length,
                      "reference_ms": (t1 - t0) * 1e3, "feed_ms": (t2 - t1) * 1e3}))
    ref.close_all()
    feed.close_all()


if __name__ == "__main__":
    main(sys.argv[1])
