"""`B200FlatIPStore`: drop-in for the reference's ``FAISSStore`` component.

Mirror of ``/root/reference/src/pipeline/components/faiss_store.py:21-189`` -- same constructor
(``settings``), ``load() / search(embeddings, k) / unload()``, ``is_loaded`` / ``index_size``
properties, ``settings`` / ``index_path`` attributes, the same exceptions with the same messages
(the reference's tests match on them, ``/root/reference/tests/test_components.py:47,92,112,133``)
and the same ``(distances, indices)`` return contract -- but the index lives in B200 HBM and
``search`` runs the hand-written sm_100a kernels behind libflatip_b200.so.  There is no CPU path.

It is looked up by the retrieval service under the registry name ``faiss_store``
(``/root/reference/src/pipeline/services/retrieval/api.py:218,383,537``); see ``factory.py``.

``settings.faiss_index_path`` may name
  * a faiss flat index file (``IxFI``) -- what the reference's ``faiss.read_index`` consumes,
  * a faiss IVF-Flat index file (``IwFl``, METRIC_L2) -- what the reference's own builder writes;
    it is served by the IVF-Flat operator (``ivf.py``) with ``nprobe = settings.faiss_nprobe``,
  * a FIPB corpus file written by ``index_io.write_fipb`` (fp32 / fp16 / bf16 rows), or
  * ``synthetic:<rows>[?dim=..&dtype=..&seed=..]`` -- a Philox corpus generated on the device.

Multi-GPU: ``config["devices"] = [0, 1, ...]`` (or ``FLATIP_B200_DEVICES=0,1,...`` / ``all``)
row-shards the corpus over those GPUs INSIDE this one component (``FlatIPGroup``, C ABI
``fip_group_*``): the reference's single call site ``registry.get("faiss_store").search(...)``
(``api.py:383-389``) then runs one host thread + stream per device, the K4 exchange + merge over
NVLink and returns the global ``(distances, indices)`` -- nothing changes for the caller.
"""

from __future__ import annotations

import gc
import logging
import os
import threading
from pathlib import Path
from urllib.parse import parse_qs

import numpy as np

from . import index_io
from .group import FlatIPGroup, calibrate_devices
from .index import FlatIPIndex, torch_stream_handle
from .synthetic import CORPUS_SEED

logger = logging.getLogger(__name__)

_UPLOAD_CHUNK_BYTES = 256 << 20


def _parse_devices(spec) -> list[int] | None:
    """None / "" -> single device; "all" -> every visible GPU; "0,1,2" or a list -> those."""
    if spec is None or spec == "":
        return None
    if isinstance(spec, str):
        if spec.strip().lower() == "all":
            import torch  # only to count devices

            return list(range(torch.cuda.device_count()))
        spec = [p for p in spec.replace(" ", "").split(",") if p]
    devs = [int(d) for d in spec]
    return devs or None


class B200FlatIPStore:
    """Exact flat inner-product store on one or several B200s of this process (same surface as the
    reference FAISSStore)."""

    def __init__(self, settings, config: dict | None = None) -> None:
        config = dict(config or {})
        self.settings = settings
        self._raw_path = str(settings.faiss_index_path)
        self.index_path = Path(self._raw_path)
        self._index: FlatIPIndex | None = None
        self._is_loaded = False
        self._lock = threading.Lock()
        self.device = int(config.get("device", os.environ.get("FLATIP_B200_DEVICE", 0)))
        # several devices = one corpus row-sharded over them behind this one component
        self.devices = _parse_devices(config.get("devices", os.environ.get("FLATIP_B200_DEVICES")))
        if self.devices:
            self.device = self.devices[0]
        # shard sizes: "equal" (default), "auto" (proportional to the throughput every device
        # sustains, measured at load() on all of them at once) or an explicit list of weights
        self.shard_weights = config.get("shard_weights", os.environ.get("FLATIP_B200_SHARD_WEIGHTS", "equal"))
        self.used_shard_weights = None
        # None = keep the dtype the file holds (fp32 for faiss files, like the reference).
        self.storage_dtype = config.get("storage_dtype", os.environ.get("FLATIP_B200_DTYPE"))
        self.id_base = int(config.get("id_base", 0))
        self.variant = config.get("variant", os.environ.get("FLATIP_B200_VARIANT", "auto"))
        # fp32 indexes (flat and IVF) filter on an fp16 copy of the rows (+50 % device memory, exact
        # fp32 results either way); False keeps only the fp32 rows.
        self.filter_copy = str(config.get("filter_copy", os.environ.get("FLATIP_B200_FILTER_COPY", "1"))
                               ).lower() not in ("0", "false", "no")

    # ------------------------------------------------------------------------------- load
    def _new_index(self, dim: int, dtype: str, nrows: int, id_base: int):
        if self.devices and len(self.devices) > 1:
            weights = self.shard_weights
            if isinstance(weights, str):
                if weights == "auto":
                    weights = calibrate_devices(
                        self.devices, dim, dtype, batch=int(getattr(self.settings, "retrieval_batch_size", 32) or 32),
                        k=int(getattr(self.settings, "retrieval_k", 10) or 10))
                elif weights == "equal":
                    weights = None
                else:
                    weights = [float(w) for w in weights.split(",")]
            self.used_shard_weights = weights
            return FlatIPGroup(dim, dtype, self.devices, nrows, id_base=id_base, weights=weights)
        return FlatIPIndex(dim, dtype, self.device, id_base=id_base, capacity_rows=nrows)

    def _load_synthetic(self):
        spec = self._raw_path[len("synthetic:"):]
        rows_s, _, query = spec.partition("?")
        opts = {k: v[0] for k, v in parse_qs(query).items()}
        dim = int(opts.get("dim", self.settings.faiss_dim))
        dtype = opts.get("dtype", self.storage_dtype or "f32")
        seed = int(opts.get("seed", CORPUS_SEED))
        row0 = int(opts.get("row0", self.id_base))
        nrows = int(float(rows_s))
        index = self._new_index(dim, dtype, nrows, row0)
        index.add_synthetic(nrows, seed, row0)
        return index

    def _load_ivf(self):
        """A faiss IndexIVFFlat file -- the index type the reference's own builder writes
        (``/root/reference/scripts/create_test_docs.py:83-104``).  Centroids and inverted lists are
        imported unchanged; ``nprobe`` follows ``settings.faiss_nprobe`` like the reference
        (``faiss_store.py:84-93``), falling back to the value stored in the file."""
        from .ivf import IVFFlatL2Index

        f = index_io.read_faiss_ivfflat(str(self.index_path))
        if f.metric != "l2":
            raise ValueError("only METRIC_L2 IVF-Flat indexes are supported")
        nprobe = int(getattr(self.settings, "faiss_nprobe", 0) or f.nprobe or 1)
        index = IVFFlatL2Index(f.dim, f.nlist, self.device, nprobe=nprobe)
        index.set_centroids(f.centroids)
        list_numbers = np.repeat(np.arange(f.nlist, dtype=np.int32), f.sizes)
        step = max(1, _UPLOAD_CHUNK_BYTES // (f.dim * 4))
        for r0 in range(0, f.ntotal, step):
            index.add_preassigned(f.codes[r0:r0 + step], f.ids[r0:r0 + step], list_numbers[r0:r0 + step])
        logger.info("Set FAISS nprobe to %d for IVF index (searching %.2f%% of clusters)", nprobe,
                    nprobe / f.nlist * 100)
        return index

    def _load_file(self):
        if index_io.is_faiss_ivfflat(str(self.index_path)):
            return self._load_ivf()
        info = index_io.read_header(str(self.index_path))
        if info.metric != "ip":
            raise ValueError("only inner-product flat indexes are supported (file holds an "
                             f"{info.metric.upper()} index)")
        dtype = self.storage_dtype or info.dtype
        rows = index_io.open_rows(str(self.index_path), info)
        index = self._new_index(info.dim, dtype, info.ntotal, info.id_base + self.id_base)
        # Rows go to the device(s) in the width the file holds: a 16-bit corpus crosses PCIe as 16
        # bits (no fp32 widening on the host), each device pulling its own slice of the mapped file
        # through its own pinned, double-buffered staging (fip_add_host / fip_group_add_host).
        esize = 4 if info.dtype == "f32" else 2
        same_width = info.dtype == dtype or "f32" in (info.dtype, dtype)
        ndev = len(self.devices) if self.devices and len(self.devices) > 1 else 1
        step = max(1, ndev * _UPLOAD_CHUNK_BYTES // (info.dim * esize))
        for r0 in range(0, info.ntotal, step):
            chunk = rows[r0:r0 + step]
            if same_width:
                index.add_raw(chunk, info.dtype)
            else:   # fp16 file into a bf16 index or vice versa: through fp32
                if info.dtype == "bf16":
                    chunk = (np.asarray(chunk).astype(np.uint32) << np.uint32(16)).view(np.float32)
                index.add(np.asarray(chunk, dtype=np.float32))
        return index

    def load(self) -> None:
        """Bring the corpus into HBM (reference: ``FAISSStore.load``, faiss_store.py:40-111).

        Raises FileNotFoundError if the index file does not exist, RuntimeError for any other
        failure -- including a missing CUDA library or GPU (no CPU fallback).
        """
        if self._is_loaded:
            logger.info("FAISS index already loaded")
            return
        synthetic = self._raw_path.startswith("synthetic:")
        if not synthetic and not self.index_path.exists():
            msg = f"FAISS index not found at {self.index_path}"
            raise FileNotFoundError(msg)
        logger.info("Loading FAISS index from %s", self._raw_path)
        try:
            index = self._load_synthetic() if synthetic else self._load_file()
            if self.variant != "auto" and hasattr(index, "set_variant"):
                index.set_variant(self.variant)
            if not self.filter_copy:
                if hasattr(index, "set_variant"):
                    index.set_option("filter_copy", 0)
                else:
                    index.set_option("tensor_format", 1)
            self._index = index
            self._is_loaded = True
            logger.info("FAISS index loaded successfully: %d vectors, dimension=%d (B200 %s, "
                        "device%s %s)", index.ntotal, index.d, index.dtype,
                        "s" if isinstance(index, FlatIPGroup) else "",
                        self.devices if isinstance(index, FlatIPGroup) else self.device)
            if index.ntotal > 0:
                logger.info("Warming up FAISS index...")
                index.search(np.zeros((1, index.d), dtype=np.float32), 1)
                logger.info("FAISS index warmup complete")
        except Exception as e:
            logger.exception("Failed to load FAISS index")
            self._index = None
            self._is_loaded = False
            msg = f"FAISS index loading failed: {e}"
            raise RuntimeError(msg) from e

    # ------------------------------------------------------------------------------ search
    def search(self, embeddings: np.ndarray, k: int) -> tuple[np.ndarray, np.ndarray]:
        """Exact inner-product top-k (reference: ``FAISSStore.search``, faiss_store.py:113-158).

        embeddings: (batch, dim) float array.  Returns (distances float32 [batch, k], indices
        int64 [batch, k]), each row best-first; unfilled slots are -FLT_MAX / -1 as in faiss.
        """
        if not self._is_loaded or self._index is None:
            msg = "FAISS index not loaded. Call load() first."
            raise RuntimeError(msg)
        if embeddings.ndim != 2:
            msg = f"Embeddings must be 2D array, got shape {embeddings.shape}"
            raise ValueError(msg)
        if embeddings.shape[1] != self.settings.faiss_dim:
            msg = (
                f"Embedding dimension mismatch: expected {self.settings.faiss_dim}, "
                f"got {embeddings.shape[1]}"
            )
            raise ValueError(msg)
        self._check_k(k)
        embeddings = embeddings.astype("float32")
        try:
            return self._index.search(embeddings, k)
        except Exception:
            logger.exception("B200 flat search failed (%d queries, k=%d)", embeddings.shape[0], k)
            raise

    MAX_K = 2048   # capacity of the select kernel's shared-memory lists (csrc/flatip_api.cu kMaxK)

    def _check_k(self, k) -> None:
        """faiss accepts any k; this library stops at 2048.  A request beyond that is the caller's
        mistake (``retrieval_k`` is an unbounded setting, config/__init__.py:236), so it is a
        ``ValueError`` -- which the retrieval service maps to a client error -- rather than a
        ``FlatIPError`` from the native layer (-> HTTP 500, api.py:263-277)."""
        if not isinstance(k, (int, np.integer)) or isinstance(k, bool) or k < 1 or k > self.MAX_K:
            msg = f"k must be an integer in [1, {self.MAX_K}], got {k!r}"
            raise ValueError(msg)

    def search_lists(self, embeddings: np.ndarray, k: int) -> tuple[list[list[int]], list[list[float]]]:
        """``search`` with the result already in the shape the retrieval service builds right after
        it: ``[row.tolist() for row in indices], [row.tolist() for row in distances]``
        (/root/reference/src/pipeline/services/retrieval/api.py:390, and :413-414 on the cached
        path), as (doc_ids, distances) -- the order the caller uses.  A convenience, not a speed-up:
        measured at batch 1024, k = 100 one ``tolist()`` per matrix takes 12.7 ms against 10.2 ms
        for the per-row loop on the build container's CPU -- the cost is creating 204 800 Python
        objects, whichever loop does it.  What removes it is not converting at all: ``id_only``
        responses and ``CandidateFeed.fetch`` (candidate_feed.py) take the int64 matrix as it is."""
        distances, indices = self.search(embeddings, k)
        return indices.tolist(), distances.tolist()

    def search_device(self, embeddings, k: int):
        """Embedding hand-off fast path: a co-located embedder passes its output as a
        torch CUDA tensor ``[batch, dim]`` on this store's device (any float dtype) and gets torch
        CUDA tensors back -- no ``list[float]`` -> ``np.array`` -> ``astype`` copies
        (``/root/reference/src/pipeline/services/retrieval/api.py:69-84``, ``faiss_store.py:147``)
        and no host round trip.  Same validation, errors and (distances, indices) contract as
        ``search``; the call is ordered on the tensor's current stream and blocks until done."""
        import torch

        if not self._is_loaded or self._index is None:
            msg = "FAISS index not loaded. Call load() first."
            raise RuntimeError(msg)
        if embeddings.dim() != 2:
            msg = f"Embeddings must be 2D array, got shape {tuple(embeddings.shape)}"
            raise ValueError(msg)
        if embeddings.shape[1] != self.settings.faiss_dim:
            msg = (
                f"Embedding dimension mismatch: expected {self.settings.faiss_dim}, "
                f"got {embeddings.shape[1]}"
            )
            raise ValueError(msg)
        if not embeddings.is_cuda or embeddings.device.index != self.device:
            msg = f"search_device needs a CUDA tensor on device {self.device}"
            raise ValueError(msg)
        self._check_k(k)
        if isinstance(self._index, FlatIPGroup):
            # several devices: the queries have to reach every GPU anyway -- go through the host path
            d, i = self._index.search(embeddings.to(torch.float32).cpu().numpy(), int(k))
            return torch.from_numpy(d).to(embeddings.device), torch.from_numpy(i).to(embeddings.device)
        q = embeddings.to(torch.float32).contiguous()
        batch = q.shape[0]
        distances = torch.empty((batch, int(k)), dtype=torch.float32, device=q.device)
        indices = torch.empty((batch, int(k)), dtype=torch.int64, device=q.device)
        if not hasattr(self._index, "set_variant"):   # IVF runs on its own stream
            torch.cuda.current_stream(q.device).synchronize()
        try:
            self._index.search_device(q.data_ptr(), batch, int(k), distances.data_ptr(),
                                      indices.data_ptr(),
                                      torch_stream_handle(q.device))
        except Exception:
            logger.exception("FAISS search failed")
            raise
        return distances, indices

    # ----------------------------------------------------------------------------- teardown
    def unload(self) -> None:
        if self._is_loaded:
            logger.info("Unloading FAISS index")
            if self._index is not None:
                self._index.close()
            self._index = None
            self._is_loaded = False
            gc.collect()

    @property
    def is_loaded(self) -> bool:
        return self._is_loaded

    @property
    def index_size(self) -> int:
        if not self._is_loaded or self._index is None:
            return 0
        return int(self._index.ntotal)

    @property
    def native_index(self):
        return self._index

    def __repr__(self) -> str:
        status = "loaded" if self._is_loaded else "not loaded"
        size = self.index_size if self._is_loaded else "unknown"
        return f"B200FlatIPStore(path={self._raw_path}, status={status}, size={size})"
