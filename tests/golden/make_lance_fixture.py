"""Extracts the reference-written IVF_PQ fixtures used by tests/test_reader.py.

Runs only where /root/reference exists (this container).  It packs, byte for byte, the legacy
`index.idx` files the reference keeps as test data (Apache-2.0, like the rest of the reference) into
`lance_files/reference_index_files.npz` —
  test_data/v0.8.14/corrupt_index/_indices/*/index.idx          (Lance 0.8.14; dim 128, IVF4, PQ16; 1000 / 2000 rows)
  test_data/v0.21.0/bad_index_fragment_bitmap/_indices/*/index.idx (Lance 0.21.0; dim 16, IVF1, PQ1; 256 + 32-row delta)
(used by rust/lance/src/dataset.rs:4262,4392) — and pulls the indexed column's raw vectors out of the
datasets' data files into `raw_vectors.npz`, keyed by row id = (fragment id << 32) | row offset, so that the
tests can check every stored PQ code and partition against the oracle without any Lance code.

Data-file parsing (test tooling only):
  * Lance v1 data file (0.8.14): footer → `pb::Metadata{manifest_position, batch_offsets, page_table_position}`
    (protos/file.proto), page table = i64 (position, length) per (field, batch)
    (rust/lance-file/src/page_table.rs:54-96); a FixedSizeList<f32> page is `rows × dim` plain f32.
  * Lance v2 data file (0.21.0): 40-byte footer (column-metadata start, CMO table, GBO table, counts, version),
    `ColumnMetadata.pages[].buffer_offsets/buffer_sizes` (protos/file2.proto); the flat FSL<f32> page is buffer 0.
  * manifest: `pb::Manifest.fragments[] {id, files[] {path}, physical_rows}` (protos/table.proto).
"""
import os
import struct
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
from lance_b200.reader import _fields, _packed_varints, read_footer  # noqa: E402

REF = "/root/reference/test_data"
OUT = os.path.join(os.path.dirname(__file__), "lance_files")


def framed(blob, pos):
    (n,) = struct.unpack_from("<I", blob, pos)
    return blob[pos + 4:pos + 4 + n]


def manifest_fragments(path):
    blob = open(path, "rb").read()
    pos, _, _ = read_footer(blob)
    frags = {}
    for fno, wt, v in _fields(framed(blob, pos)):
        if fno == 2:
            fid, files = 0, []
            for f2, w2, v2 in _fields(v):
                if f2 == 1:
                    fid = v2
                elif f2 == 2:
                    for f3, w3, v3 in _fields(v2):
                        if f3 == 1:
                            files.append(bytes(v3).decode())
            frags[fid] = files[0]
    return frags


def read_v1_vectors(path, dim, field_id=0, num_fields=4):
    blob = open(path, "rb").read()
    pos, major, minor = read_footer(blob)
    assert (major, minor) == (0, 1), (major, minor)
    batch_offsets, pt_pos = [], None
    for fno, wt, v in _fields(framed(blob, pos)):
        if fno == 2:
            _packed_varints(v, wt, batch_offsets)
        elif fno == 3:
            pt_pos = v
    nb = len(batch_offsets) - 1
    table = np.frombuffer(blob, dtype="<i8", count=num_fields * nb * 2, offset=pt_pos).reshape(num_fields, nb, 2)
    out = []
    for b in range(nb):
        p, n = int(table[field_id, b, 0]), int(table[field_id, b, 1])
        assert n == batch_offsets[b + 1] - batch_offsets[b]
        out.append(np.frombuffer(blob, dtype="<f4", count=n * dim, offset=p).reshape(n, dim))
    return np.concatenate(out)


def read_v2_vectors(path, dim):
    blob = open(path, "rb").read()
    assert blob[-4:] == b"LANC"
    cm_start, cmo, gbo, ngb, ncol, major, minor = struct.unpack("<QQQIIHH", blob[-40:-4])
    assert ncol == 1 and (major, minor) == (0, 3), (ncol, major, minor)
    cpos, clen = struct.unpack_from("<QQ", blob, cmo)
    out = []
    for fno, wt, v in _fields(blob[cpos:cpos + clen]):
        if fno == 2:   # Page
            offs, sizes, length = [], [], 0
            for f2, w2, v2 in _fields(v):
                if f2 == 1:
                    _packed_varints(v2, w2, offs)
                elif f2 == 2:
                    _packed_varints(v2, w2, sizes)
                elif f2 == 3:
                    length = v2
            assert sizes[0] == length * dim * 4, (sizes, length)
            out.append(np.frombuffer(blob, dtype="<f4", count=length * dim, offset=offs[0]).reshape(length, dim))
    return np.concatenate(out)


def main():
    os.makedirs(OUT, exist_ok=True)
    # ---- v0.8.14 -------------------------------------------------------------------------------------
    src = f"{REF}/v0.8.14/corrupt_index"
    frags = {}
    for ver in sorted(os.listdir(f"{src}/_versions")):
        frags.update(manifest_fragments(f"{src}/_versions/{ver}"))
    rid, vec = [], []
    for fid, fn in sorted(frags.items()):
        x = read_v1_vectors(f"{src}/data/{fn}", 128)
        rid.append((np.uint64(fid) << np.uint64(32)) | np.arange(len(x), dtype=np.uint64))
        vec.append(x)
    np.savez_compressed(f"{OUT}/v0.8.14_raw_vectors.npz", row_ids=np.concatenate(rid), vectors=np.concatenate(vec))
    packed = {}
    for u in os.listdir(f"{src}/_indices"):
        packed[f"v0.8.14_{u[:8]}"] = np.frombuffer(open(f"{src}/_indices/{u}/index.idx", "rb").read(), dtype=np.uint8)
    # ---- v0.21.0 -------------------------------------------------------------------------------------
    src = f"{REF}/v0.21.0/bad_index_fragment_bitmap"
    frags = {}
    for ver in sorted(os.listdir(f"{src}/_versions")):
        frags.update(manifest_fragments(f"{src}/_versions/{ver}"))
    rid, vec = [], []
    for fid, fn in sorted(frags.items()):
        x = read_v2_vectors(f"{src}/data/{fn}", 16)
        rid.append((np.uint64(fid) << np.uint64(32)) | np.arange(len(x), dtype=np.uint64))
        vec.append(x)
    np.savez_compressed(f"{OUT}/v0.21.0_raw_vectors.npz", row_ids=np.concatenate(rid), vectors=np.concatenate(vec))
    for u in os.listdir(f"{src}/_indices"):
        packed[f"v0.21.0_{u[:8]}"] = np.frombuffer(open(f"{src}/_indices/{u}/index.idx", "rb").read(), dtype=np.uint8)
    # the index files' bytes, unchanged, in one archive (tests/conftest.py::lance_files_dir writes them back out)
    np.savez_compressed(f"{OUT}/reference_index_files.npz", **packed)
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(f"{OUT}/{f}"))


if __name__ == "__main__":
    main()
