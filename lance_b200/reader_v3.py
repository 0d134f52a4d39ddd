"""Reader for the CURRENT on-disk IVF_PQ index of Lance (`IndexFileVersion::V3`: `_indices/<uuid>/index.idx` +
`auxiliary.idx`, both Lance-v2 file containers) -> the exchange format `lb200_index_create` takes.

PARITY UNPINNED: the reference tree holds no v3 index file (`find /root/reference -name auxiliary.idx` is empty) and
the Rust writer cannot run in this image, so this reader is pinned only on files produced by `tests/lance_v3_writer.py`,
a restatement of the writer below from the same sources.  What it follows:

  * file container (protos/file2.proto, rust/lance-file/src/v2/writer.rs:395-590): data buffers, column metadatas, the
    column-metadata offset table (pos u64, len u64 per column), the global-buffer offset table (pos u64, len u64), and
    the 40-byte footer `[col-meta start u64][CMO table u64][GBO table u64][#global buffers u32][#columns u32][major u16]
    [minor u16]"LANC"`; version (0,3) or (2,0) = LanceFileVersion::V2_0 (lance-encoding/src/version.rs:38-60).
    Global buffer 0 = `pb::FileDescriptor{schema{fields, metadata}, length}` (writer.rs:413-440, protos/file.proto).
  * page encodings of format 2.0 (protos/encodings.proto; lance-encoding/src/encoder.rs:519-660): u64 ->
    `Nullable{no_nulls{Flat{bits_per_value 64, buffer}}}`; FixedSizeList<u8, M> ->
    `Nullable{no_nulls{FixedSizeList{dimension M, items Nullable{no_nulls{Flat{8}}}}}}`, wrapped in `google.protobuf.Any`
    inside `Encoding.direct` (writer.rs:178-206).  Anything else (compression, bit packing, nulls) is rejected.
  * what the vector index writes (rust/lance/src/index/vector/builder.rs:624-770): index.idx carries the schema metadata
    `lance:index` = {"type": "IVF_PQ", "distance_type": ...} and `lance:ivf` = the global-buffer index of a `pb::Ivf`
    with the centroids tensor; auxiliary.idx carries `distance_type`, `lance:ivf` (a `pb::Ivf` with lengths only) and
    `storage_metadata` = a JSON list with ONE JSON string: ProductQuantizationMetadata {num_bits, num_sub_vectors,
    dimension, codebook_tensor (pb::Tensor bytes as a JSON number array), transposed}
    (rust/lance-index/src/vector/pq.rs:438-457, pq/storage.rs:52-65), and the columns `_rowid` (u64) and `__pq_code`
    (FixedSizeList<u8, M>).  With `transposed` every partition's slice of the `__pq_code` value buffer is already
    `[M][n_p]` (pq/storage.rs:455-470), i.e. the exchange format; otherwise it is transposed here (pq/storage.rs:357-377).
  * how it is read back: IVFIndex::try_new (rust/lance/src/index/vector/ivf/v2.rs:110-202),
    IvfQuantizationStorage::try_new / load_partition (rust/lance-index/src/vector/storage.rs:210-298).
"""
import json
import struct

import numpy as np

from .reader import MAGIC, _fields, _ivf, _packed_varints, _tensor
from .synth import IvfPqData

FOOTER_LEN = 40
_ANY_ARRAY_ENCODING = "/lance.encodings.ArrayEncoding"


def _sub(buf, want):
    """The length-delimited sub-message in field `want` of `buf`, or None."""
    for fno, wt, v in _fields(buf):
        if fno == want and wt == 2:
            return v
    return None


class V2File:
    """The parts of a Lance-v2 file this reader needs: schema fields + metadata, global buffers, flat column values."""

    def __init__(self, blob, what):
        self.blob = blob = memoryview(bytes(blob))
        self.what = what
        n = len(blob)
        if n < FOOTER_LEN:
            raise ValueError(f"{what}: does not have sufficient data, len: {n}")
        if bytes(blob[n - 4:]) != MAGIC:
            raise ValueError(f"{what}: not a Lance file: magic number mismatch")
        cm_start, cmo, gbo, ngb, ncol, major, minor = struct.unpack_from("<QQQIIHH", blob, n - FOOTER_LEN)
        if (major, minor) not in ((0, 3), (2, 0)):
            raise ValueError(f"{what}: file version {major}.{minor} is not the Lance 2.0 container of a v3 vector index")
        end = n - FOOTER_LEN
        if not (cm_start <= cmo <= gbo <= end) or (gbo - cmo) != 16 * ncol or (end - gbo) != 16 * ngb:
            raise ValueError(f"{what}: inconsistent footer offsets")
        self.columns_meta = [self._slice(*struct.unpack_from("<QQ", blob, cmo + 16 * i), "column metadata") for i in range(ncol)]
        self.global_buffers = [struct.unpack_from("<QQ", blob, gbo + 16 * i) for i in range(ngb)]
        if ngb < 1:
            raise ValueError(f"{what}: no global buffers (the schema lives in buffer 0)")
        self._parse_descriptor(self.global_buffer(0))

    def _slice(self, pos, length, name):
        if pos > len(self.blob) or length > len(self.blob) - pos:
            raise ValueError(f"{self.what}: {name} runs past the end of the file")
        return self.blob[pos:pos + length]

    def global_buffer(self, index):
        if index >= len(self.global_buffers):
            raise ValueError(f"{self.what}: global buffer {index} does not exist")
        return self._slice(*self.global_buffers[index], "global buffer")

    def _parse_descriptor(self, buf):
        self.fields, self.metadata, self.num_rows = [], {}, 0      # fields: (name, id, parent_id, logical_type)
        for fno, wt, v in _fields(buf):
            if fno == 1 and wt == 2:                               # Schema
                for f2, w2, v2 in _fields(v):
                    if f2 == 1 and w2 == 2:                        # Field
                        name, fid, parent, ltype = "", 0, 0, ""
                        for f3, w3, v3 in _fields(v2):
                            if f3 == 2 and w3 == 2:
                                name = bytes(v3).decode()
                            elif f3 == 3:
                                fid = v3 - (1 << 64) if v3 >> 63 else v3
                            elif f3 == 4:
                                parent = v3 - (1 << 64) if v3 >> 63 else v3
                            elif f3 == 5 and w3 == 2:
                                ltype = bytes(v3).decode()
                        self.fields.append((name, fid, parent, ltype))
                    elif f2 == 5 and w2 == 2:                      # map<string, bytes> entry
                        k, val = b"", b""
                        for f3, w3, v3 in _fields(v2):
                            if f3 == 1 and w3 == 2:
                                k = bytes(v3)
                            elif f3 == 2 and w3 == 2:
                                val = bytes(v3)
                        self.metadata[k.decode()] = val.decode()
            elif fno == 2:
                self.num_rows = v

    def column_index(self, name):
        """Top-level primitive / fixed-size-list fields take one column each, in schema order (format 2.0)."""
        tops = [f for f in self.fields if f[2] < 0]                # parent_id -1 = top level (datatypes/field.rs:931)
        for i, f in enumerate(tops):
            if f[0] == name:
                return i, f[3]
        raise ValueError(f"{self.what}: column {name} not found")

    @staticmethod
    def _flat(enc, bits):
        """ArrayEncoding -> its Flat leaf, through Nullable{no_nulls}; checks bits per value and no compression."""
        nullable = _sub(enc, 2)
        if nullable is not None:
            if _sub(nullable, 2) is not None or _sub(nullable, 3) is not None:
                raise ValueError("null values are not supported in index columns")
            enc = _sub(_sub(nullable, 1), 1)                       # NoNull.values
        flat = _sub(enc, 1)
        if flat is None:
            raise ValueError("unsupported page encoding (expected Flat)")
        got, bufidx, compressed = 0, 0, False
        for fno, wt, v in _fields(flat):
            if fno == 1:
                got = v
            elif fno == 2 and wt == 2:
                for f2, _, v2 in _fields(v):
                    if f2 == 1:
                        bufidx = v2
            elif fno == 3:
                compressed = True
        if got != bits or compressed:
            raise ValueError(f"unsupported Flat page encoding (bits {got}, compressed {compressed})")
        return bufidx

    def column_values(self, col, bytes_per_row, fsl_dim=None):
        """All pages of column `col` concatenated: `rows * bytes_per_row` bytes."""
        if col >= len(self.columns_meta):
            raise ValueError(f"{self.what}: column {col} has no metadata")
        parts, rows = [], 0
        for fno, wt, page in _fields(self.columns_meta[col]):
            if fno != 2 or wt != 2:
                continue
            offs, sizes, length, enc = [], [], 0, None
            for f2, w2, v2 in _fields(page):
                if f2 == 1:
                    _packed_varints(v2, w2, offs)
                elif f2 == 2:
                    _packed_varints(v2, w2, sizes)
                elif f2 == 3:
                    length = v2
                elif f2 == 4 and w2 == 2:
                    enc = v2
            direct = _sub(enc, 2) if enc is not None else None     # Encoding.direct
            any_msg = _sub(direct, 1) if direct is not None else None
            if any_msg is None:
                raise ValueError(f"{self.what}: page without a direct encoding")
            url, payload = "", None
            for f2, w2, v2 in _fields(any_msg):
                if f2 == 1 and w2 == 2:
                    url = bytes(v2).decode()
                elif f2 == 2 and w2 == 2:
                    payload = v2
            if not url.endswith(_ANY_ARRAY_ENCODING.lstrip("/")) or payload is None:
                raise ValueError(f"{self.what}: page encoding {url!r} is not a format-2.0 ArrayEncoding")
            if fsl_dim is not None:
                nullable = _sub(payload, 2)
                inner = _sub(_sub(nullable, 1), 1) if nullable is not None else payload
                fsl = _sub(inner, 3)
                if fsl is None:
                    raise ValueError(f"{self.what}: expected a FixedSizeList page encoding")
                dim, items = 0, None
                for f2, w2, v2 in _fields(fsl):
                    if f2 == 1:
                        dim = v2
                    elif f2 == 2 and w2 == 2:
                        items = v2
                if dim != fsl_dim or items is None:
                    raise ValueError(f"{self.what}: FixedSizeList dimension {dim} != {fsl_dim}")
                bufidx = self._flat(items, 8)
            else:
                bufidx = self._flat(payload, 8 * bytes_per_row)
            if bufidx >= len(offs) or len(offs) != len(sizes):
                raise ValueError(f"{self.what}: page buffer {bufidx} missing")
            need = length * bytes_per_row
            if sizes[bufidx] < need:
                raise ValueError(f"{self.what}: page buffer holds {sizes[bufidx]} bytes, {need} needed")
            parts.append(self._slice(offs[bufidx], need, "page buffer"))
            rows += length
        out = np.frombuffer(b"".join(bytes(p) for p in parts), dtype=np.uint8) if parts else np.zeros(0, np.uint8)
        return out, rows


def open_v3_ivf_pq(index_idx, auxiliary_idx) -> IvfPqData:
    """(index.idx, auxiliary.idx) as paths or bytes -> IvfPqData (codes transposed per partition).  Malformed input
    raises ValueError."""
    try:
        return _open_v3(index_idx, auxiliary_idx)
    except (TypeError, KeyError, IndexError, OverflowError, struct.error, UnicodeDecodeError, AttributeError) as e:
        raise ValueError(f"malformed v3 index files: {type(e).__name__}: {e}") from None


def _open_v3(index_idx, auxiliary_idx) -> IvfPqData:
    def load(x):
        return open(x, "rb").read() if isinstance(x, str) else bytes(x)
    idx = V2File(load(index_idx), "index.idx")
    aux = V2File(load(auxiliary_idx), "auxiliary.idx")
    try:
        meta = json.loads(idx.metadata["lance:index"])
        index_type, metric = meta["type"], meta["distance_type"].lower()
    except KeyError as e:
        raise ValueError(f"index.idx: schema metadata {e} not found") from None
    if index_type != "IVF_PQ":
        raise ValueError(f"index type {index_type} is not IVF_PQ")
    if metric not in ("l2", "cosine", "dot"):
        raise ValueError(f"Metric type '{metric}' is not supported")
    if aux.metadata.get("distance_type", metric).lower() != metric:
        raise ValueError("index.idx and auxiliary.idx disagree on the distance type")
    ivf = _ivf(idx.global_buffer(int(idx.metadata["lance:ivf"])))
    sivf = _ivf(aux.global_buffer(int(aux.metadata["lance:ivf"])))
    if ivf.centroids is None:
        raise ValueError("index.idx: IVF model has no centroids")
    nlist, dim = ivf.centroids.shape
    lengths = np.asarray(sivf.lengths, dtype=np.int64)
    if len(lengths) != nlist:
        raise ValueError(f"auxiliary.idx: {len(lengths)} partitions, index.idx has {nlist}")
    pq = json.loads(json.loads(aux.metadata["storage_metadata"])[0])
    M, bits = int(pq["num_sub_vectors"]), int(pq["num_bits"])
    if bits != 8:
        raise ValueError(f"num_bits {bits}: only 8-bit PQ is on the B200 path")
    if int(pq["dimension"]) != dim or dim % M:
        raise ValueError("PQ metadata does not match the IVF dimension")
    codebook = _tensor(bytes(bytearray(pq["codebook_tensor"])))
    if codebook.size != 256 * dim:
        raise ValueError(f"PQ codebook has {codebook.size} values, expected 256x{dim}")
    N = int(lengths.sum())
    rid_col, _ = aux.column_index("_rowid")
    pq_col, _ = aux.column_index("__pq_code")
    rid_bytes, n1 = aux.column_values(rid_col, 8)
    code_bytes, n2 = aux.column_values(pq_col, M, fsl_dim=M)
    if n1 != N or n2 != N:
        raise ValueError(f"auxiliary.idx holds {n1}/{n2} rows, the IVF model {N}")
    offs = np.zeros(nlist + 1, dtype=np.uint64)
    offs[1:] = np.cumsum(lengths)
    codes = code_bytes.copy()
    if not pq.get("transposed", False):                            # row-major per partition -> [M][n_p]
        for p in range(nlist):
            o, e = int(offs[p]), int(offs[p + 1])
            codes[o * M:e * M] = code_bytes[o * M:e * M].reshape(e - o, M).T.reshape(-1)
    return IvfPqData(metric, dim, nlist, M, np.ascontiguousarray(ivf.centroids, dtype=np.float32),
                     np.ascontiguousarray(codebook.reshape(M, 256, dim // M), dtype=np.float32), offs, codes,
                     np.frombuffer(rid_bytes.tobytes(), dtype="<u8").astype(np.uint64))
