"""TEST TOOL: restatement of how the reference WRITES a v3 IVF_PQ index (`index.idx` + `auxiliary.idx`), used to produce
files for the v3 reader (lance_b200/reader_v3.py, csrc/api_file_v3.inl) -- the reference tree holds no such fixture and
the Rust writer cannot run here, so the v3 reader is PARITY UNPINNED and says so.

Follows rust/lance/src/index/vector/builder.rs:624-770 (what goes into the two files), rust/lance-file/src/v2/writer.rs
(page buffers padded to 64 bytes with 0x48, pages in batch order, column metadatas, CMO / GBO tables, 40-byte footer,
version (0,3) for format 2.0), protos/file2.proto + encodings.proto (page / column messages) and
rust/lance-index/src/vector/pq.rs:438-457 (ProductQuantizationMetadata as JSON, codebook tensor as a number array)."""
import json
import struct

import numpy as np

PAD = 64


def _vi(x):
    x &= (1 << 64) - 1
    out = bytearray()
    while True:
        b = x & 0x7F
        x >>= 7
        out.append(b | (0x80 if x else 0))
        if not x:
            return bytes(out)


def _ld(fno, payload):
    return _vi(fno << 3 | 2) + _vi(len(payload)) + bytes(payload)


def _vf(fno, x):
    return _vi(fno << 3) + _vi(x)


def tensor(a):
    a = np.ascontiguousarray(a, dtype="<f4")
    return _vf(1, 2) + _ld(2, b"".join(_vi(s) for s in a.shape)) + _ld(3, a.tobytes())


def _flat(bits, bufidx=0):
    return _ld(1, _vf(1, bits) + _ld(2, _vf(1, bufidx)))                     # ArrayEncoding.flat{bits, buffer{index, page}}


def _no_nulls(values):
    return _ld(2, _ld(1, _ld(1, values)))                                     # ArrayEncoding.nullable{no_nulls{values}}


def _any(msg):
    return _ld(1, b"type.googleapis.com/lance.encodings.ArrayEncoding") + _ld(2, msg)


def _field(name, fid, ltype, nullable=False):
    return (_vf(1, 2) + _ld(2, name.encode()) + _vf(3, fid) + _vf(4, -1) + _ld(5, ltype.encode()) + _vf(6, int(nullable)))


class V2Writer:
    def __init__(self, fields):
        self.buf = bytearray()
        self.fields = fields                 # [(name, logical type)]
        self.pages = [[] for _ in fields]    # per column: encoded pb Page messages
        self.global_buffers = []
        self.metadata = {}
        self.rows = 0

    def _write_buffer(self, b):
        pos = len(self.buf)
        self.buf += b
        self.buf += b"\x48" * ((-len(b)) % PAD)
        return pos

    def write_page(self, col, data, rows, encoding, priority):
        pos = self._write_buffer(data)
        enc = _ld(2, _ld(1, _any(encoding)))                                  # Encoding.direct{encoding = Any}
        self.pages[col].append(_ld(1, _vi(pos)) + _ld(2, _vi(len(data))) + _vf(3, rows) + _ld(4, enc) + _vf(5, priority))

    def add_global_buffer(self, b):
        pos = self._write_buffer(b)
        self.global_buffers.append((pos, len(b)))
        return len(self.global_buffers)                                       # indices start at 1 (0 = the schema)

    def finish(self, major=0, minor=3):
        schema = b"".join(_ld(1, _field(n, i, t)) for i, (n, t) in enumerate(self.fields))
        for k, v in self.metadata.items():
            schema += _ld(5, _ld(1, k.encode()) + _ld(2, v.encode()))
        desc = _ld(1, schema) + _vf(2, self.rows)
        gbo = [(len(self.buf), len(desc))]
        self.buf += desc
        gbo += self.global_buffers
        cm_start = len(self.buf)
        cm_pos = []
        for pages in self.pages:
            values_enc = _ld(1, b"type.googleapis.com/lance.encodings.ColumnEncoding") + _ld(2, _ld(1, b""))
            meta = _ld(1, _ld(2, _ld(1, values_enc))) + b"".join(_ld(2, p) for p in pages)
            cm_pos.append((len(self.buf), len(meta)))
            self.buf += meta
        cmo = len(self.buf)
        for pos, n in cm_pos:
            self.buf += struct.pack("<QQ", pos, n)
        gbo_start = len(self.buf)
        for pos, n in gbo:
            self.buf += struct.pack("<QQ", pos, n)
        self.buf += struct.pack("<QQQIIHH", cm_start, cmo, gbo_start, len(gbo), len(self.fields), major, minor) + b"LANC"
        return bytes(self.buf)


def write_v3(data, transposed=True, rows_per_page=None, version=(0, 3), num_bits=8):
    """IvfPqData -> (index.idx bytes, auxiliary.idx bytes)."""
    nlist, M, dim = data.nlist, data.M, data.dim
    lengths = np.diff(data.part_offsets.astype(np.int64))
    # ---- auxiliary.idx: one batch per non-empty partition (builder.rs:657-683) ----
    aux = V2Writer([("_rowid", "uint64"), ("__pq_code", f"fixed_size_list:uint8:{M}")])
    row = 0
    for p in range(nlist):
        o, e = int(data.part_offsets[p]), int(data.part_offsets[p + 1])
        if e == o:
            continue
        rids = data.row_ids[o:e].astype("<u8").tobytes()
        codes_t = data.codes[o * M:e * M].reshape(M, e - o)
        codes = (codes_t if transposed else codes_t.T).reshape(-1).tobytes()   # the FSL value buffer of the batch
        step = rows_per_page or (e - o)
        for s in range(0, e - o, step):                                         # a batch may be cut into several pages
            n = min(step, e - o - s)
            aux.write_page(0, rids[s * 8:(s + n) * 8], n, _no_nulls(_flat(64)), row + s)
            fsl = _ld(3, _vf(1, M) + _ld(2, _no_nulls(_flat(8))))              # fixed_size_list{dimension, items}
            aux.write_page(1, codes[s * M:(s + n) * M], n, _no_nulls(fsl), row + s)
        row += e - o
    aux.rows = row
    aux.metadata["distance_type"] = data.metric
    sivf = _ld(3, b"".join(_vi(int(n)) for n in lengths))                       # storage IVF: lengths only (+ offsets)
    sivf = _ld(2, b"".join(_vi(int(x)) for x in data.part_offsets[:-1])) + sivf
    aux.metadata["lance:ivf"] = str(aux.add_global_buffer(sivf))
    pqm = {"codebook_position": 0, "num_bits": num_bits, "num_sub_vectors": M, "dimension": dim,
           "codebook_tensor": list(tensor(data.codebook.reshape(-1, dim // M))), "transposed": bool(transposed)}
    aux.metadata["storage_metadata"] = json.dumps([json.dumps(pqm, separators=(",", ":"))])
    # ---- index.idx: FLAT sub-index has no rows; the IVF model with the centroids (builder.rs:685-760) ----
    idx = V2Writer([("__flat_marker", "uint64")])
    idx.metadata["lance:index"] = json.dumps({"type": "IVF_PQ", "distance_type": data.metric})
    iivf = _ld(2, b"".join(_vi(0) for _ in range(nlist))) + _ld(3, b"".join(_vi(0) for _ in range(nlist))) + _ld(4, tensor(data.centroids))
    idx.metadata["lance:ivf"] = str(idx.add_global_buffer(iivf))
    idx.metadata["lance:flat"] = json.dumps([""] * nlist)
    return idx.finish(*version), aux.finish(*version)
