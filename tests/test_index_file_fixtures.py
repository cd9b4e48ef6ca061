"""index_io's readers against byte-level fixtures written by an INDEPENDENT writer
(tests/golden/make_index_fixtures.py: numpy records laid down from the documented faiss layout, no
import of index_io), so a mistake shared by index_io's reader and its own writer cannot hide.
Files: what faiss.write_index emits for IndexFlatIP / IndexFlatL2 / IndexIVFFlat
(/root/reference/src/pipeline/components/faiss_store.py:66-69,
/root/reference/scripts/create_test_docs.py:83-104,
/root/reference/tests/test_retrieval_service.py:205-212)."""
import os

import numpy as np
import pytest

from rag_inference_pipeline_b200 import index_io

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
EXPECTED = np.load(os.path.join(GOLDEN, "index_file_fixtures_expected.npz"))


@pytest.mark.parametrize("name,metric", [("flat_ip_5x8.IxFI.bin", "ip"), ("flat_l2_5x8.IxF2.bin", "l2")])
def test_flat_reader_on_independent_bytes(name, metric):
    path = os.path.join(GOLDEN, name)
    info = index_io.read_header(path)
    assert (info.fmt, info.dim, info.ntotal, info.dtype, info.metric) == ("faiss_flat", 8, 5, "f32", metric)
    assert info.data_offset == 4 + 33 + 8 and os.path.getsize(path) == info.data_offset + 5 * 8 * 4
    assert np.array_equal(np.asarray(index_io.open_rows(path, info)), EXPECTED["rows"])


def test_flat_reader_on_the_reference_fixture_recipe():
    path = os.path.join(GOLDEN, "reference_fixture_flat_l2_10x768.IxF2.bin")
    info = index_io.read_header(path)
    assert (info.dim, info.ntotal, info.metric) == (768, 10, "l2")
    want = np.random.default_rng(42).random((10, 768)).astype("float32")
    assert np.array_equal(np.asarray(index_io.open_rows(path, info)), want)


def test_flat_writer_emits_the_same_bytes_as_the_independent_writer(tmp_path):
    for name, metric in (("flat_ip_5x8.IxFI.bin", "ip"), ("flat_l2_5x8.IxF2.bin", "l2")):
        out = tmp_path / name
        index_io.write_faiss_flat(str(out), EXPECTED["rows"], metric)
        assert out.read_bytes() == open(os.path.join(GOLDEN, name), "rb").read()


def test_ivfflat_reader_full_size_table():
    path = os.path.join(GOLDEN, "ivfflat_l2_4lists_6x8_full.IwFl.bin")
    assert index_io.is_faiss_ivfflat(path)
    f = index_io.read_faiss_ivfflat(path)
    assert (f.dim, f.ntotal, f.nlist, f.nprobe, f.metric) == (8, 6, 4, 2, "l2")
    assert np.array_equal(f.centroids, EXPECTED["centroids"])
    assert np.array_equal(f.sizes, EXPECTED["sizes"])
    assert np.array_equal(f.codes, EXPECTED["codes"]) and np.array_equal(f.ids, EXPECTED["ids"])


def test_ivfflat_reader_sparse_size_table():
    f = index_io.read_faiss_ivfflat(os.path.join(GOLDEN, "ivfflat_l2_8lists_3x8_sprs.IwFl.bin"))
    assert (f.dim, f.ntotal, f.nlist, f.nprobe) == (8, 3, 8, 1)
    assert f.sizes.tolist() == [1, 0, 0, 0, 0, 0, 0, 2]
    assert np.array_equal(f.centroids, EXPECTED["cent8"])
    assert np.array_equal(f.codes, EXPECTED["codes"][:3]) and np.array_equal(f.ids, EXPECTED["ids"][:3])


def test_ivfflat_writer_emits_the_same_bytes_as_the_independent_writer(tmp_path):
    out = tmp_path / "w.bin"
    index_io.write_faiss_ivfflat(str(out), EXPECTED["centroids"], EXPECTED["sizes"], EXPECTED["codes"],
                                 EXPECTED["ids"], nprobe=2, metric="l2")
    assert out.read_bytes() == open(os.path.join(GOLDEN, "ivfflat_l2_4lists_6x8_full.IwFl.bin"), "rb").read()


def test_truncated_and_inconsistent_files_are_refused(tmp_path):
    raw = open(os.path.join(GOLDEN, "flat_ip_5x8.IxFI.bin"), "rb").read()
    short = tmp_path / "short.bin"
    short.write_bytes(raw[:-4])
    with pytest.raises(ValueError, match="shorter than its header"):
        index_io.read_header(str(short))
    bad = bytearray(raw)
    bad[4 + 33] ^= 1                                   # vector word count no longer ntotal * d
    wrong = tmp_path / "wrong.bin"
    wrong.write_bytes(bytes(bad))
    with pytest.raises(ValueError, match="inconsistent"):
        index_io.read_header(str(wrong))
