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


# ------------------------------------------------------------------------------------------
# FAISS IVF-Flat index files: what the reference's own builder writes
# (/root/reference/scripts/create_test_docs.py:83-104, `faiss.write_index(IndexIVFFlat)`) and
# FAISSStore.load reads (faiss_store.py:66-69).  Layout restated from upstream faiss
# `impl/index_write.cpp` (write_index / write_ivf_header / write_direct_map / write_InvertedLists);
# like the flat layout above it could not be checked against a real faiss file in this image, so
# the reader validates every field it can and fails loudly on anything unexpected:
#   "IwFl", index header (d:int32, ntotal:int64, dummy:int64, dummy:int64, is_trained:uint8,
#   metric_type:int32 [+ metric_arg:float32 when metric_type > 1]), nlist:uint64, nprobe:uint64,
#   the quantizer as a nested flat index ("IxF2" / "IxFI" + header + count:uint64 + fp32 rows),
#   direct map (type:uint8, array count:uint64 + int64 entries [+ hashtable pairs when type == 2]),
#   inverted lists "ilar": nlist:uint64, code_size:uint64 (= 4 d), list sizes as "full"
#   (count:uint64 + uint64 sizes) or "sprs" (count:uint64 + (list, size) uint64 pairs), then for
#   every non-empty list its codes (size * code_size bytes) followed by its ids (size int64).
FAISS_FOURCC_IVFFLAT = b"IwFl"


@dataclass
class IVFFlatFile:
    dim: int
    ntotal: int
    nlist: int
    nprobe: int
    metric: str              # "l2" | "ip"
    centroids: np.ndarray    # [nlist, dim] fp32
    sizes: np.ndarray        # [nlist] int64
    codes: np.ndarray        # [ntotal, dim] fp32, concatenated in list order
    ids: np.ndarray          # [ntotal] int64


def is_faiss_ivfflat(path: str) -> bool:
    with open(path, "rb") as f:
        return f.read(4) == FAISS_FOURCC_IVFFLAT


def _read_index_header(f):
    fixed = f.read(33)
    if len(fixed) != 33:
        raise ValueError("truncated faiss index header")
    d, ntotal, _d1, _d2, trained, metric_type = struct.unpack("<iqqqBi", fixed)
    if metric_type > 1:
        f.read(4)
    if d <= 0 or ntotal < 0 or metric_type not in (METRIC_INNER_PRODUCT, METRIC_L2):
        raise ValueError(f"unsupported faiss index header: d={d} ntotal={ntotal} metric={metric_type}")
    return d, ntotal, bool(trained), "ip" if metric_type == METRIC_INNER_PRODUCT else "l2"


def _read_exact(f, n: int) -> bytes:
    b = f.read(n)
    if len(b) != n:
        raise ValueError("faiss index file is shorter than its header claims")
    return b


def read_faiss_ivfflat(path: str) -> IVFFlatFile:
    with open(path, "rb") as f:
        if f.read(4) != FAISS_FOURCC_IVFFLAT:
            raise ValueError("not a faiss IndexIVFFlat file (fourcc IwFl)")
        d, ntotal, _trained, metric = _read_index_header(f)
        nlist, nprobe = struct.unpack("<QQ", _read_exact(f, 16))
        if nlist < 1 or nlist > (1 << 24):
            raise ValueError(f"implausible nlist {nlist}")
        qcc = f.read(4)
        if qcc not in (FAISS_FOURCC_IP, FAISS_FOURCC_L2):
            raise ValueError(f"unsupported coarse quantizer {qcc!r} (expected a flat index)")
        qd, qn, _qt, _qm = _read_index_header(f)
        (count,) = struct.unpack("<Q", _read_exact(f, 8))
        if qd != d or qn != nlist or count != nlist * d:
            raise ValueError(f"inconsistent coarse quantizer: d={qd} ntotal={qn} words={count}")
        centroids = np.frombuffer(_read_exact(f, count * 4), dtype=np.float32).reshape(nlist, d).copy()
        (dm_type,) = struct.unpack("<B", _read_exact(f, 1))
        (dm_n,) = struct.unpack("<Q", _read_exact(f, 8))
        f.seek(dm_n * 8, 1)
        if dm_type == 2:
            (hn,) = struct.unpack("<Q", _read_exact(f, 8))
            f.seek(hn * 16, 1)
        elif dm_type not in (0, 1):
            raise ValueError(f"unknown direct map type {dm_type}")
        if _read_exact(f, 4) != b"ilar":
            raise ValueError("unsupported inverted lists (expected ArrayInvertedLists 'ilar')")
        il_nlist, code_size = struct.unpack("<QQ", _read_exact(f, 16))
        if il_nlist != nlist or code_size != 4 * d:
            raise ValueError(f"inconsistent inverted lists: nlist={il_nlist} code_size={code_size}")
        kind = _read_exact(f, 4)
        sizes = np.zeros(nlist, dtype=np.int64)
        (cnt,) = struct.unpack("<Q", _read_exact(f, 8))
        if kind == b"full":
            if cnt != nlist:
                raise ValueError("inverted list size table has the wrong length")
            sizes[:] = np.frombuffer(_read_exact(f, 8 * nlist), dtype=np.uint64).astype(np.int64)
        elif kind == b"sprs":
            pairs = np.frombuffer(_read_exact(f, 8 * cnt), dtype=np.uint64).astype(np.int64).reshape(-1, 2)
            sizes[pairs[:, 0]] = pairs[:, 1]
        else:
            raise ValueError(f"unknown inverted list size encoding {kind!r}")
        if int(sizes.sum()) != ntotal:
            raise ValueError(f"inverted lists hold {int(sizes.sum())} vectors, header says {ntotal}")
        codes = np.empty((ntotal, d), dtype=np.float32)
        ids = np.empty(ntotal, dtype=np.int64)
        pos = 0
        for l in range(nlist):
            n = int(sizes[l])
            if n == 0:
                continue
            codes[pos:pos + n] = np.frombuffer(_read_exact(f, n * code_size), dtype=np.float32).reshape(n, d)
            ids[pos:pos + n] = np.frombuffer(_read_exact(f, n * 8), dtype=np.int64)
            pos += n
    return IVFFlatFile(d, ntotal, int(nlist), int(nprobe), metric, centroids, sizes, codes, ids)


def write_faiss_ivfflat(path: str, centroids: np.ndarray, sizes: np.ndarray, codes: np.ndarray,
                        ids: np.ndarray, nprobe: int = 1, metric: str = "l2") -> None:
    centroids = np.ascontiguousarray(centroids, dtype=np.float32)
    codes = np.ascontiguousarray(codes, dtype=np.float32)
    ids = np.ascontiguousarray(ids, dtype=np.int64)
    sizes = np.ascontiguousarray(sizes, dtype=np.int64)
    nlist, d = centroids.shape
    ntotal = int(sizes.sum())
    if codes.shape != (ntotal, d) or ids.shape != (ntotal,) or sizes.shape != (nlist,):
        raise ValueError("inconsistent inverted lists")
    mt = METRIC_INNER_PRODUCT if metric == "ip" else METRIC_L2
    with open(path, "wb") as f:
        f.write(FAISS_FOURCC_IVFFLAT)
        f.write(struct.pack("<iqqqBi", d, ntotal, 1 << 20, 1 << 20, 1, mt))
        f.write(struct.pack("<QQ", nlist, nprobe))
        f.write(FAISS_FOURCC_L2 if metric == "l2" else FAISS_FOURCC_IP)
        f.write(struct.pack("<iqqqBi", d, nlist, 1 << 20, 1 << 20, 1, mt))
        f.write(struct.pack("<Q", nlist * d))
        f.write(centroids.tobytes())
        f.write(struct.pack("<BQ", 0, 0))           # direct map: NoMap, empty array
        f.write(b"ilar")
        f.write(struct.pack("<QQ", nlist, 4 * d))
        nonzero = np.flatnonzero(sizes)
        if len(nonzero) > nlist // 2:
            f.write(b"full")
            f.write(struct.pack("<Q", nlist))
            f.write(sizes.astype(np.uint64).tobytes())
        else:
            f.write(b"sprs")
            f.write(struct.pack("<Q", 2 * len(nonzero)))
            f.write(np.stack([nonzero, sizes[nonzero]], axis=1).astype(np.uint64).tobytes())
        pos = 0
        for l in range(nlist):
            n = int(sizes[l])
            if n:
                f.write(codes[pos:pos + n].tobytes())
                f.write(ids[pos:pos + n].tobytes())
                pos += n
