"""CPU tests of the drop-in boundary: the C-ABI library loads and exports every symbol the header
declares, fails loudly without a GPU (no CPU fallback), and the store mirrors the reference
FAISSStore's error behaviour (/root/reference/tests/test_components.py:34-184)."""
import os
import re
import types

import numpy as np
import pytest

from conftest import ROOT, gpu_available

HEADER = os.path.join(ROOT, "include", "flatip_b200.h")


def _settings(path, dim=768):
    return types.SimpleNamespace(faiss_index_path=str(path), faiss_dim=dim, faiss_use_mmap=False)


def test_library_exports_every_declared_symbol(native_lib):
    from rag_inference_pipeline_b200 import _native

    text = open(HEADER).read()
    declared = set(re.findall(r"\b(fip_[a-z0-9_]+)\s*\(", text))
    declared -= {"fip_index"}
    assert declared == set(_native.EXPORTED_SYMBOLS), declared ^ set(_native.EXPORTED_SYMBOLS)
    for name in declared:
        assert hasattr(native_lib, name), name
    assert native_lib.fip_abi_version() == 1
    assert native_lib.fip_compiled_sm() == 100


def test_library_is_sm100a_tcgen05(native_lib):
    """The shipped cubin must carry the Blackwell-native opcodes (no legacy HMMA path)."""
    import shutil
    import subprocess

    from rag_inference_pipeline_b200 import _native

    if not shutil.which("cuobjdump"):
        pytest.skip("cuobjdump not on PATH")
    sass = subprocess.run(["cuobjdump", "-sass", _native.LIB_PATH], capture_output=True, text=True).stdout
    assert "sm_100a" in sass
    assert "UTCHMMA" in sass      # tcgen05.mma
    assert "LDTM" in sass         # tcgen05.ld
    assert "UTMALDG" in sass      # TMA
    assert " HMMA" not in sass    # no mma.sync / wmma


@pytest.mark.skipif(gpu_available(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback_without_gpu(native_lib, tmp_path):
    from rag_inference_pipeline_b200 import B200FlatIPStore, FlatIPIndex, index_io
    from rag_inference_pipeline_b200._native import FIP_ERR_NODEVICE, FlatIPError

    with pytest.raises(FlatIPError) as ei:
        FlatIPIndex(64, "f32")
    assert ei.value.code == FIP_ERR_NODEVICE
    path = tmp_path / "idx.faiss"
    index_io.write_faiss_flat(str(path), np.zeros((4, 64), np.float32))
    store = B200FlatIPStore(_settings(path, 64))
    with pytest.raises(RuntimeError, match="FAISS index loading failed"):
        store.load()
    assert not store.is_loaded and store.index_size == 0


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    from rag_inference_pipeline_b200 import _native

    monkeypatch.setattr(_native, "_lib", None)
    monkeypatch.setattr(_native, "LIB_PATH", str(tmp_path / "nope.so"))
    with pytest.raises(_native.NativeLibraryError, match="no CPU fallback"):
        _native.load_library()


# ---- mirrors of the reference's TestFAISSStore (tests/test_components.py) -------------------
def test_init(tmp_path):
    from rag_inference_pipeline_b200 import B200FlatIPStore

    s = _settings(tmp_path / "faiss_index.bin")
    store = B200FlatIPStore(s)
    assert store.settings is s
    assert str(store.index_path) == str(tmp_path / "faiss_index.bin")
    assert store._index is None
    assert store.is_loaded is False
    assert store.index_size == 0
    assert "not loaded" in repr(store)


def test_filter_copy_switch_from_config_and_env(tmp_path, monkeypatch):
    """fp32 indexes filter on an fp16 copy of the rows unless the profile / environment says no."""
    from rag_inference_pipeline_b200 import B200FlatIPStore

    s = _settings(tmp_path / "faiss_index.bin")
    assert B200FlatIPStore(s).filter_copy is True
    assert B200FlatIPStore(s, {"filter_copy": False}).filter_copy is False
    assert B200FlatIPStore(s, {"filter_copy": "0"}).filter_copy is False
    monkeypatch.setenv("FLATIP_B200_FILTER_COPY", "false")
    assert B200FlatIPStore(s).filter_copy is False
    assert B200FlatIPStore(s, {"filter_copy": True}).filter_copy is True


def test_load_file_not_found(tmp_path):
    from rag_inference_pipeline_b200 import B200FlatIPStore

    store = B200FlatIPStore(_settings(tmp_path / "missing.bin"))
    with pytest.raises(FileNotFoundError, match="FAISS index not found"):
        store.load()


def test_search_not_loaded_raises_error(tmp_path):
    from rag_inference_pipeline_b200 import B200FlatIPStore

    store = B200FlatIPStore(_settings(tmp_path / "x.bin"))
    with pytest.raises(RuntimeError, match="not loaded"):
        store.search(np.random.rand(1, 768).astype("float32"), k=10)


class _FakeIndex:
    d, ntotal, dtype = 768, 10, "f32"

    def search(self, x, k):
        n = x.shape[0]
        return np.zeros((n, k), np.float32), np.tile(np.arange(k, dtype=np.int64), (n, 1))

    def close(self):
        pass


def _loaded_store(tmp_path):
    from rag_inference_pipeline_b200 import B200FlatIPStore

    store = B200FlatIPStore(_settings(tmp_path / "x.bin"))
    store._index = _FakeIndex()
    store._is_loaded = True
    return store


def test_search_invalid_embedding_shape(tmp_path):
    store = _loaded_store(tmp_path)
    with pytest.raises(ValueError, match="2D array"):
        store.search(np.random.rand(768).astype("float32"), k=10)


def test_search_wrong_dimension(tmp_path):
    store = _loaded_store(tmp_path)
    with pytest.raises(ValueError, match="dimension mismatch"):
        store.search(np.random.rand(1, 512).astype("float32"), k=10)


def test_search_passes_through_and_casts(tmp_path):
    store = _loaded_store(tmp_path)
    d, i = store.search(np.random.rand(3, 768), k=5)  # float64 in, like np.array(list_of_lists)
    assert d.shape == (3, 5) and d.dtype == np.float32
    assert i.shape == (3, 5) and i.dtype == np.int64
    assert isinstance(i[0].tolist()[0], int) and isinstance(d[0].tolist()[0], float)


def test_search_k_outside_the_library_limit_is_a_value_error(tmp_path):
    """faiss takes any k, this library stops at 2048: the store says so with the exception type the
    retrieval service treats as a bad request, before the native layer is reached."""
    store = _loaded_store(tmp_path)
    x = np.random.rand(2, 768).astype("float32")
    for bad in (0, -1, 2049, 10.0, True):
        with pytest.raises(ValueError, match=r"k must be an integer in \[1, 2048\]"):
            store.search(x, k=bad)
    d, i = store.search(x, k=2048)
    assert i.shape == (2, 2048)
    assert store.search(x, k=np.int64(7))[1].shape == (2, 7)


def test_search_lists_equals_the_callers_per_row_loop(tmp_path):
    """api.py:390 builds ``[row.tolist() for row in indices], [row.tolist() for row in distances]``;
    search_lists returns the same nested lists (same Python types) from one tolist() per matrix."""
    store = _loaded_store(tmp_path)
    x = np.random.rand(5, 768).astype("float32")
    distances, indices = store.search(x, 6)
    ids, dists = store.search_lists(x, 6)
    assert ids == [row.tolist() for row in indices] and dists == [row.tolist() for row in distances]
    assert type(ids[0][0]) is int and type(dists[0][0]) is float


def test_unload_and_is_loaded(tmp_path):
    store = _loaded_store(tmp_path)
    assert store.is_loaded and store.index_size == 10
    store.unload()
    assert not store.is_loaded and store._index is None and store.index_size == 0


def test_factory_and_install():
    from rag_inference_pipeline_b200 import B200FlatIPStore, COMPONENT_TYPE, create_b200_flat_ip_store, install

    s = _settings("/nonexistent")
    assert isinstance(create_b200_flat_ip_store(s, {}), B200FlatIPStore)
    factories = {"faiss_store": object(), "document_store": "keep"}
    install(factories)
    assert factories["faiss_store"] is create_b200_flat_ip_store
    assert factories[COMPONENT_TYPE] is create_b200_flat_ip_store
    assert factories["document_store"] == "keep"


# ---- index file formats ---------------------------------------------------------------------
def test_faiss_flat_file_roundtrip(tmp_path):
    from rag_inference_pipeline_b200 import index_io

    rows = np.random.default_rng(0).standard_normal((37, 24)).astype(np.float32)
    p = str(tmp_path / "flat.index")
    index_io.write_faiss_flat(p, rows)
    info = index_io.read_header(p)
    assert (info.fmt, info.dim, info.ntotal, info.dtype, info.metric) == ("faiss_flat", 24, 37, "f32", "ip")
    assert np.array_equal(np.asarray(index_io.open_rows(p, info)), rows)
    raw = open(p, "rb").read()
    assert raw[:4] == b"IxFI" and len(raw) == 4 + 33 + 8 + rows.nbytes
    index_io.write_faiss_flat(p, rows, metric="l2")
    assert index_io.read_header(p).metric == "l2"
    with open(p, "r+b") as f:
        f.truncate(100)
    with pytest.raises(ValueError):
        index_io.read_header(p)


@pytest.mark.parametrize("dtype", ["f32", "f16", "bf16"])
def test_fipb_roundtrip(tmp_path, dtype):
    from rag_inference_pipeline_b200 import index_io
    from rag_inference_pipeline_b200.synthetic import round_to_storage, bf16_bits_to_f32

    rows = np.random.default_rng(1).standard_normal((19, 40)).astype(np.float32)
    p = str(tmp_path / "c.fipb")
    index_io.write_fipb(p, rows, dtype, id_base=1000)
    info = index_io.read_header(p)
    assert (info.fmt, info.dim, info.ntotal, info.dtype, info.id_base) == ("fipb", 40, 19, dtype, 1000)
    got = np.asarray(index_io.open_rows(p, info))
    got = bf16_bits_to_f32(got) if dtype == "bf16" else got.astype(np.float32)
    assert np.array_equal(got, round_to_storage(rows, dtype))


def test_unrecognised_file(tmp_path):
    from rag_inference_pipeline_b200 import index_io

    p = tmp_path / "junk.bin"
    p.write_bytes(b"IwFl" + b"\0" * 100)   # an IVF fourcc: not a flat index
    with pytest.raises(ValueError, match="unrecognised"):
        index_io.read_header(str(p))


def test_group_abi_fails_loudly_without_a_gpu(native_lib, tmp_path):
    """The multi-device entry points behave like the single-device ones on a box without a GPU:
    argument errors are reported as such, everything else is FIP_ERR_NODEVICE -- never a CPU path."""
    import ctypes

    from rag_inference_pipeline_b200 import B200FlatIPStore, _native
    from rag_inference_pipeline_b200.group import FlatIPGroup, calibrate_devices

    if gpu_available():
        pytest.skip("this check is for boxes without a GPU")
    cfg = _native.FipConfig(dim=64, dtype=1, device=0, reserved=0, capacity_rows=0, id_base=0)
    devs = (ctypes.c_int32 * 2)(0, 1)
    h = ctypes.c_void_p()
    assert native_lib.fip_group_create(ctypes.byref(cfg), devs, 0, 100, ctypes.byref(h)) == _native.FIP_ERR_INVALID
    assert native_lib.fip_group_create(ctypes.byref(cfg), devs, 2, -1, ctypes.byref(h)) == _native.FIP_ERR_INVALID
    w = (ctypes.c_double * 2)(1.0, 0.0)
    assert native_lib.fip_group_create_weighted(ctypes.byref(cfg), devs, 2, 100, w, ctypes.byref(h)) in (
        _native.FIP_ERR_INVALID, _native.FIP_ERR_NODEVICE)
    assert native_lib.fip_group_create(ctypes.byref(cfg), devs, 2, 100, ctypes.byref(h)) == _native.FIP_ERR_NODEVICE
    assert not h.value
    assert native_lib.fip_group_search(None, None, 1, 1, None, None) == _native.FIP_ERR_INVALID
    flagged = ctypes.c_int32(7)
    assert native_lib.fip_search_status(None, ctypes.byref(flagged)) == _native.FIP_ERR_INVALID
    with pytest.raises(_native.FlatIPError) as e:
        FlatIPGroup(64, "f16", [0, 1], 1000)
    assert e.value.code == _native.FIP_ERR_NODEVICE
    with pytest.raises(_native.FlatIPError):
        calibrate_devices([0], 64, "f16", rows=10)           # bad argument (rows < 1024)
    with pytest.raises(_native.FlatIPError):
        calibrate_devices([0], 64, "f16", rows=4096, seconds=0.01)   # no device
    store = B200FlatIPStore(_settings("synthetic:1000?dim=768&dtype=f16"), {"devices": [0, 1]})
    with pytest.raises(RuntimeError, match="FAISS index loading failed"):
        store.load()
    assert not store.is_loaded


def test_store_device_list_parsing():
    from rag_inference_pipeline_b200.store import _parse_devices

    assert _parse_devices(None) is None and _parse_devices("") is None
    assert _parse_devices("0,1, 2") == [0, 1, 2] and _parse_devices([3, 1]) == [3, 1]
    assert _parse_devices("5") == [5]
