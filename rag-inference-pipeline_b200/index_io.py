"""On-disk corpus formats for the flat-IP store.

* FAISS flat index files (what ``faiss.write_index(IndexFlatIP)`` emits and the reference loads
  with ``faiss.read_index``, ``/root/reference/src/pipeline/components/faiss_store.py:66-69``):
  fourcc ``IxFI`` (inner product) / ``IxF2`` (L2), the generic index header
  ``d:int32, ntotal:int64, dummy:int64, dummy:int64, is_trained:uint8, metric_type:int32``
  (``+ metric_arg:float32`` when ``metric_type > 1``), then the code vector as ``count:uint64``
  (number of 32-bit words = ntotal * d) followed by the raw fp32 rows.  Layout restated from
  upstream faiss ``impl/index_write.cpp``; it could not be checked against a real faiss file in
  this image (faiss is not installable), so the reader validates every field it can.
* ``FIPB``: this package's own format -- 8-byte magic, uint32 JSON-header length, JSON header
  (dim, dtype, ntotal, id_base), padding to 64 bytes, raw rows in the storage dtype.  Row-sharded
  corpora are one FIPB file per shard.
"""

from __future__ import annotations

import json
import os
import struct
from dataclasses import dataclass

import numpy as np

FAISS_FOURCC_IP = b"IxFI"
FAISS_FOURCC_L2 = b"IxF2"
FIPB_MAGIC = b"FIPB0001"
METRIC_INNER_PRODUCT = 0
METRIC_L2 = 1


@dataclass
class CorpusFileInfo:
    fmt: str          # "faiss_flat" | "fipb"
    dim: int
    ntotal: int
    dtype: str        # dtype of the rows in the file: "f32" | "f16" | "bf16"
    metric: str       # "ip" | "l2"
    data_offset: int
    id_base: int = 0


_NP_FILE_DTYPE = {"f32": np.float32, "f16": np.float16, "bf16": np.uint16}


def read_header(path: str) -> CorpusFileInfo:
    size = os.path.getsize(path)
    with open(path, "rb") as f:
        head = f.read(8)
        if head[:4] in (FAISS_FOURCC_IP, FAISS_FOURCC_L2):
            f.seek(4)
            fixed = f.read(4 + 8 + 8 + 8 + 1 + 4)
            if len(fixed) != 33:
                raise ValueError("truncated faiss index header")
            d, ntotal, _d1, _d2, _trained, metric_type = struct.unpack("<iqqqBi", fixed)
            if metric_type > 1:
                f.read(4)
            (count,) = struct.unpack("<Q", f.read(8))
            if d <= 0 or ntotal < 0 or count != ntotal * d:
                raise ValueError(f"inconsistent faiss flat index header: d={d} ntotal={ntotal} "
                                 f"vector words={count}")
            off = f.tell()
            if size < off + count * 4:
                raise ValueError("faiss flat index file is shorter than its header claims")
            metric = "ip" if head[:4] == FAISS_FOURCC_IP else "l2"
            return CorpusFileInfo("faiss_flat", d, ntotal, "f32", metric, off)
        if head == FIPB_MAGIC:
            (hlen,) = struct.unpack("<I", f.read(4))
            meta = json.loads(f.read(hlen).decode("utf-8"))
            off = (12 + hlen + 63) // 64 * 64
            info = CorpusFileInfo("fipb", int(meta["dim"]), int(meta["ntotal"]), meta["dtype"],
                                  meta.get("metric", "ip"), off, int(meta.get("id_base", 0)))
            esize = 4 if info.dtype == "f32" else 2
            if size < off + info.ntotal * info.dim * esize:
                raise ValueError("FIPB file is shorter than its header claims")
            return info
    raise ValueError(f"unrecognised index file (fourcc {head[:4]!r}); expected a faiss flat index "
                     "(IxFI/IxF2) or a FIPB corpus")


def open_rows(path: str, info: CorpusFileInfo) -> np.ndarray:
    """Memory-map the rows of a corpus file ([ntotal, dim] in the file's dtype)."""
    if info.ntotal == 0:
        return np.zeros((0, info.dim), dtype=_NP_FILE_DTYPE[info.dtype])
    return np.memmap(path, dtype=_NP_FILE_DTYPE[info.dtype], mode="r", offset=info.data_offset,
                     shape=(info.ntotal, info.dim))


def write_faiss_flat(path: str, rows: np.ndarray, metric: str = "ip") -> None:
    rows = np.ascontiguousarray(rows, dtype=np.float32)
    n, d = rows.shape
    with open(path, "wb") as f:
        f.write(FAISS_FOURCC_IP if metric == "ip" else FAISS_FOURCC_L2)
        f.write(struct.pack("<iqqqBi", d, n, 1 << 20, 1 << 20, 1,
                            METRIC_INNER_PRODUCT if metric == "ip" else METRIC_L2))
        f.write(struct.pack("<Q", n * d))
        f.write(rows.tobytes())


def write_fipb(path: str, rows: np.ndarray, dtype: str = "f32", id_base: int = 0) -> None:
    """rows: fp32 values [n, dim]; stored rounded (RNE) to `dtype`."""
    from .synthetic import round_to_bf16_bits

    rows = np.ascontiguousarray(rows, dtype=np.float32)
    n, d = rows.shape
    if dtype == "f32":
        payload = rows
    elif dtype == "f16":
        payload = rows.astype(np.float16)
    elif dtype == "bf16":
        payload = round_to_bf16_bits(rows)
    else:
        raise ValueError(f"unknown dtype {dtype!r}")
    meta = json.dumps({"dim": d, "ntotal": n, "dtype": dtype, "metric": "ip",
                       "id_base": id_base}).encode("utf-8")
    with open(path, "wb") as f:
        f.write(FIPB_MAGIC)
        f.write(struct.pack("<I", len(meta)))
        f.write(meta)
        f.write(b"\0" * ((12 + len(meta) + 63) // 64 * 64 - 12 - len(meta)))
        f.write(np.ascontiguousarray(payload).tobytes())
