"""Writes the byte-level index-file fixtures under tests/golden/ (run from the repo root:
python tests/golden/make_index_fixtures.py).

An INDEPENDENT writer: it imports nothing from the package and lays the bytes down field by field
from the documented faiss serialisation (last bullet; upstream
impl/index_write.cpp: write_index_header, WRITEXBVECTOR, write_ivf_header, write_direct_map,
write_InvertedLists / ArrayInvertedLists), with numpy structured records instead of struct.pack,
so that a mistake shared by index_io's reader AND its own writer shows up.  The files are what the
reference would hand to faiss.read_index (/root/reference/src/pipeline/components/faiss_store.py:66-69;
/root/reference/scripts/create_test_docs.py:83-104 writes the IVF one).  PARITY UNPINNED all the same:
no real faiss file exists in this image to compare these bytes with.
"""
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))

# write_index_header: int d; idx_t ntotal; idx_t dummy = 1 << 20 (twice); bool is_trained; int metric_type
INDEX_HEADER = np.dtype([("d", "<i4"), ("ntotal", "<i8"), ("dummy0", "<i8"), ("dummy1", "<i8"),
                         ("is_trained", "u1"), ("metric_type", "<i4")])   # packed: 33 bytes
assert INDEX_HEADER.itemsize == 33


def index_header(d, ntotal, metric_type):
    h = np.zeros(1, dtype=INDEX_HEADER)
    h["d"], h["ntotal"], h["dummy0"], h["dummy1"] = d, ntotal, 1 << 20, 1 << 20
    h["is_trained"], h["metric_type"] = 1, metric_type
    return h.tobytes()


def u64(*values):
    return np.asarray(values, dtype="<u8").tobytes()


def flat_index(rows, l2):
    """IndexFlatIP ("IxFI", metric 0) / IndexFlatL2 ("IxF2", metric 1): header, then the code vector as
    a count of 4-byte words followed by the fp32 rows."""
    n, d = rows.shape
    return (b"IxF2" if l2 else b"IxFI") + index_header(d, n, 1 if l2 else 0) + u64(n * d) + rows.astype("<f4").tobytes()


def ivf_flat_index(centroids, lists, nprobe, sparse_sizes):
    """IndexIVFFlat, METRIC_L2, IndexFlatL2 quantizer, no direct map, ArrayInvertedLists.
    lists: [(codes [n_l, d] fp32, ids [n_l] int64)] per inverted list."""
    nlist, d = centroids.shape
    ntotal = sum(len(ids) for _, ids in lists)
    out = [b"IwFl", index_header(d, ntotal, 1), u64(nlist, nprobe)]
    out.append(flat_index(centroids, l2=True))                     # the quantizer, a whole index
    out += [np.asarray([0], dtype="u1").tobytes(), u64(0)]          # direct map: type NoMap, empty array
    out += [b"ilar", u64(nlist, 4 * d)]
    sizes = [len(ids) for _, ids in lists]
    if sparse_sizes:
        pairs = [v for l, n in enumerate(sizes) if n for v in (l, n)]
        out += [b"sprs", u64(len(pairs)), u64(*pairs)]
    else:
        out += [b"full", u64(nlist), u64(*sizes)]
    for codes, ids in lists:
        if len(ids):
            out += [codes.astype("<f4").tobytes(), np.asarray(ids, dtype="<i8").tobytes()]
    return b"".join(out)


def main():
    rng = np.random.default_rng(20240607)
    rows = rng.standard_normal((5, 8)).astype(np.float32)
    with open(os.path.join(HERE, "flat_ip_5x8.IxFI.bin"), "wb") as f:
        f.write(flat_index(rows, l2=False))
    with open(os.path.join(HERE, "flat_l2_5x8.IxF2.bin"), "wb") as f:
        f.write(flat_index(rows, l2=True))
    # the reference's only real-index fixture: IndexFlatL2 over default_rng(42).random((10, 768))
    # (/root/reference/tests/test_retrieval_service.py:205-212)
    ref = np.random.default_rng(42).random((10, 768)).astype("float32")
    with open(os.path.join(HERE, "reference_fixture_flat_l2_10x768.IxF2.bin"), "wb") as f:
        f.write(flat_index(ref, l2=True))
    centroids = rng.standard_normal((4, 8)).astype(np.float32)
    sizes = [3, 0, 2, 1]
    ids = rng.permutation(6).astype(np.int64)
    codes = rng.standard_normal((6, 8)).astype(np.float32)
    lists, pos = [], 0
    for n in sizes:
        lists.append((codes[pos:pos + n], ids[pos:pos + n]))
        pos += n
    with open(os.path.join(HERE, "ivfflat_l2_4lists_6x8_full.IwFl.bin"), "wb") as f:
        f.write(ivf_flat_index(centroids, lists, nprobe=2, sparse_sizes=False))
    sparse = [(codes[:1], ids[:1])] + [(codes[:0], ids[:0])] * 6 + [(codes[1:3], ids[1:3])]
    cent8 = rng.standard_normal((8, 8)).astype(np.float32)
    with open(os.path.join(HERE, "ivfflat_l2_8lists_3x8_sprs.IwFl.bin"), "wb") as f:
        f.write(ivf_flat_index(cent8, sparse, nprobe=1, sparse_sizes=True))
    np.savez(os.path.join(HERE, "index_file_fixtures_expected.npz"), rows=rows, centroids=centroids,
             sizes=np.asarray(sizes, dtype=np.int64), ids=ids, codes=codes, cent8=cent8)


if __name__ == "__main__":
    main()
