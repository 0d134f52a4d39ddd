"""CPU tests of lance_b200/reader.py against index files WRITTEN BY THE REFERENCE (Lance 0.8.14 and 0.21.0,
kept by the reference as test data: rust/lance/src/dataset.rs:4262,4392) and against files produced by a
minimal writer of the same legacy layout (`write_pq_partitions`, rust/lance/src/index/vector/ivf/io.rs; metadata
`pb::Index`, protos/index.proto).

The strongest check here pins the ORACLE as well as the reader: every PQ code and every partition assignment the
reference wrote (48 288 codes over four files) equals what oracle.pq_quantize / oracle.find_partitions compute
from the dataset's raw vectors, the file's centroids and the file's codebook.
"""
import hashlib
import os
import struct

import numpy as np
import pytest

from lance_b200.reader import open_legacy_ivf_pq, parse_index_proto, read_footer
from lance_b200.synth import IvfPqData

from conftest import lance_files_dir  # noqa: E402

G = lance_files_dir()
FILES = {
    # tag, uuid prefix: (name, dataset_version, dim, nlist, M, lengths)
    ("v0.8.14", "d32dac97"): ("vector_idx", 1, 128, 4, 16, [251, 236, 249, 264]),
    ("v0.8.14", "f358a219"): ("vector_idx", 4, 128, 4, 16, [516, 471, 501, 512]),
    ("v0.21.0", "dc833a6e"): ("vector_idx", 1, 16, 1, 1, [256]),
    ("v0.21.0", "ca9b1111"): ("_vector_idx", 3, 16, 1, 1, [32]),
}


def _path(tag, u):
    return os.path.join(G, f"{tag}_{u}_index.idx")


def _raw(tag):
    z = np.load(os.path.join(G, f"{tag}_raw_vectors.npz"))
    lut = {int(r): i for i, r in enumerate(z["row_ids"].tolist())}
    vec = z["vectors"]
    return lambda ids: vec[[lut[int(r)] for r in np.asarray(ids).tolist()]]


@pytest.mark.parametrize("key", sorted(FILES))
def test_reads_reference_written_index(key):
    name, version, dim, nlist, M, lengths = FILES[key]
    li = open_legacy_ivf_pq(_path(*key))
    m, d = li.meta, li.data
    assert (m.name, m.columns, m.dataset_version, m.spec_version) == (name, ["vector"], version, 1)
    assert (m.dimension, m.metric, m.pq.num_bits) == (dim, "l2", 8)
    assert (d.dim, d.nlist, d.M, d.N) == (dim, nlist, M, sum(lengths))
    assert m.ivf.lengths == lengths and d.part_offsets.tolist() == [0] + np.cumsum(lengths).tolist()
    assert d.centroids.shape == (nlist, dim) and d.codebook.shape == (M, 256, dim // M)
    assert d.codes.shape == (d.N * M,) and d.row_ids.shape == (d.N,)
    assert len(np.unique(d.row_ids)) == d.N                      # every row indexed once
    assert np.isfinite(d.centroids).all() and np.isfinite(d.codebook).all()
    # row id = (fragment id << 32) | row offset (rust/lance-core/src/utils/address.rs)
    assert int((d.row_ids & np.uint64(0xFFFFFFFF)).max()) < 1000 and int((d.row_ids >> np.uint64(32)).max()) <= 10
    # partition pages are laid back to back from byte 0 and end where the metadata starts
    blob = open(_path(*key), "rb").read()
    pos, major, minor = read_footer(blob)
    assert (major, minor) == (0, 1)
    exp = np.concatenate([[0], np.cumsum([n * (M + 8) for n in lengths])])
    assert m.ivf.offsets == exp[:-1].tolist() and pos == exp[-1]


@pytest.mark.parametrize("key", sorted(FILES))
def test_reference_written_codes_equal_oracle_quantisation(oracle, key):
    """Pins reader + oracle on reference output: partition = argmin centroid (kmeans.rs:624-644 with nprobes 1),
    code = per-sub-space argmin over the codebook of the RESIDUAL (pq.rs:98-180; residual for L2:
    pq/builder.rs:66-68, ivf/v2.rs:285-301)."""
    li = open_legacy_ivf_pq(_path(*key))
    d = li.data
    raw = _raw(key[0])(d.row_ids)
    part = np.repeat(np.arange(d.nlist), np.diff(d.part_offsets.astype(np.int64)))
    got = np.array([oracle.find_partitions(d.centroids, x, 1, "l2")[0][0] for x in raw])
    assert got.tolist() == part.tolist()
    codes = oracle.pq_quantize("l2", d.codebook, raw - d.centroids[part])       # [N][M]
    stored = np.concatenate([d.part_codes(p).T for p in range(d.nlist)])        # back to row-major [N][M]
    assert codes.shape == stored.shape
    assert (codes != stored).sum() == 0


def test_reference_written_index_self_query(oracle):
    """Domain property on the real index: searching with an indexed row's own vector, all partitions probed and
    refine on the raw vectors, returns that row first with distance exactly 0."""
    li = open_legacy_ivf_pq(_path("v0.8.14", "f358a219"), raw_lookup=_raw("v0.8.14"))
    ox = oracle.OracleIndex(li.data)
    pick = np.arange(0, li.data.N, 97)
    od, oi, oc, _ = ox.search(li.data.raw_vectors[pick], 1, 4, refine_factor=20)
    assert oc.tolist() == [1] * len(pick)
    assert od[:, 0].tolist() == [0.0] * len(pick)
    # duplicates of a vector may exist; the winner is then the smallest row id among them
    for j, p in enumerate(pick):
        same = np.nonzero((li.data.raw_vectors == li.data.raw_vectors[p]).all(1))[0]
        assert int(oi[j, 0]) == int(li.data.row_ids[same].min())


@pytest.mark.skipif(not os.path.isdir("/root/reference/test_data"), reason="reference tree not present on this box")
def test_committed_fixtures_are_the_reference_bytes():
    src = {"v0.8.14": "/root/reference/test_data/v0.8.14/corrupt_index/_indices",
           "v0.21.0": "/root/reference/test_data/v0.21.0/bad_index_fragment_bitmap/_indices"}
    for (tag, u) in FILES:
        full = [x for x in os.listdir(src[tag]) if x.startswith(u)][0]
        a = hashlib.sha256(open(os.path.join(src[tag], full, "index.idx"), "rb").read()).hexdigest()
        b = hashlib.sha256(open(_path(tag, u), "rb").read()).hexdigest()
        assert a == b


# ---- a minimal writer of the same layout (test tool): covers branches the reference's fixtures do not ------------
def _vi(x):
    out = bytearray()
    while True:
        b = x & 0x7F
        x >>= 7
        out.append(b | (0x80 if x else 0))
        if not x:
            return bytes(out)


def _ld(fno, payload):
    return _vi(fno << 3 | 2) + _vi(len(payload)) + payload


def _vf(fno, x):
    return _vi(fno << 3) + _vi(x)


def _tensor(a):
    return _vf(1, 2) + _ld(2, b"".join(_vi(s) for s in a.shape)) + _ld(3, a.astype("<f4").tobytes())


def write_legacy(data: IvfPqData, *, tensors=True, num_bits=8, version=(0, 1), metric_code=None, name="vector_idx"):
    body, offsets = bytearray(), []
    for p in range(data.nlist):
        o, e = int(data.part_offsets[p]), int(data.part_offsets[p + 1])
        offsets.append(len(body))
        body += np.ascontiguousarray(data.part_codes(p).T).tobytes()        # row-major codes
        body += data.row_ids[o:e].astype("<u8").tobytes()
    lengths = np.diff(data.part_offsets.astype(np.int64)).tolist()
    ivf = _ld(2, b"".join(_vi(o) for o in offsets)) + _ld(3, b"".join(_vi(n) for n in lengths))
    pq = _vf(1, num_bits) + _vf(2, data.M) + _vf(3, data.dim)
    if tensors:
        ivf += _ld(4, _tensor(data.centroids))
        pq += _ld(5, _tensor(data.codebook.reshape(-1, data.dim // data.M)))
    else:   # deprecated repeated-float fields (index format before the tensors; storage.rs:184-191, pq.rs:84-87)
        ivf = _ld(1, data.centroids.astype("<f4").tobytes()) + ivf
        pq += _ld(4, data.codebook.astype("<f4").tobytes())
    mc = {"l2": 0, "cosine": 1, "dot": 2}[data.metric] if metric_code is None else metric_code
    vec = _vf(1, 1) + _vf(2, data.dim) + _ld(3, _ld(2, ivf)) + _ld(3, _ld(3, pq)) + _vf(4, mc)
    msg = _ld(1, name.encode()) + _ld(2, b"vector") + _vf(3, 7) + _vf(4, 0) + _ld(5, vec)
    pos = len(body)
    return bytes(body) + struct.pack("<I", len(msg)) + msg + struct.pack("<QHH", pos, *version) + b"LANC"


def _synthetic(metric="cosine", d=24, nlist=5, M=3, seed=3):
    rng = np.random.default_rng(seed)
    sizes = np.array([7, 0, 33, 1, 16])
    offs = np.zeros(nlist + 1, dtype=np.uint64)
    offs[1:] = np.cumsum(sizes)
    N = int(sizes.sum())
    return IvfPqData(metric, d, nlist, M, rng.normal(size=(nlist, d)).astype(np.float32),
                     rng.normal(size=(M, 256, d // M)).astype(np.float32), offs,
                     rng.integers(0, 256, size=N * M, dtype=np.uint8),
                     rng.permutation(1 << 20)[:N].astype(np.uint64))


@pytest.mark.parametrize("tensors", [True, False])
@pytest.mark.parametrize("metric", ["l2", "cosine", "dot"])
def test_writer_reader_round_trip(metric, tensors):
    data = _synthetic(metric)
    li = open_legacy_ivf_pq(write_legacy(data, tensors=tensors))
    got = li.data
    assert (got.metric, got.dim, got.nlist, got.M) == (metric, data.dim, data.nlist, data.M)
    for f in ("centroids", "codebook", "part_offsets", "codes", "row_ids"):
        assert np.array_equal(getattr(got, f), getattr(data, f)), f
    assert li.meta.dataset_version == 7 and li.meta.ivf.lengths == [7, 0, 33, 1, 16]


def test_malformed_files_raise():
    data = _synthetic("l2")
    good = write_legacy(data)
    with pytest.raises(ValueError, match="sufficient data"):
        open_legacy_ivf_pq(good[-8:])
    with pytest.raises(ValueError, match="magic"):
        open_legacy_ivf_pq(good[:-4] + b"XXXX")
    with pytest.raises(ValueError, match="legacy IVF_PQ layout"):
        open_legacy_ivf_pq(write_legacy(data, version=(0, 3)))
    with pytest.raises(ValueError, match="8-bit"):
        open_legacy_ivf_pq(write_legacy(_four_bit(data), num_bits=4))
    with pytest.raises(ValueError, match="not supported"):
        open_legacy_ivf_pq(write_legacy(data, metric_code=3))        # Hamming
    bad_pos = good[:-16] + struct.pack("<QHH", len(good), 0, 1) + b"LANC"
    with pytest.raises(ValueError, match="too small"):
        open_legacy_ivf_pq(bad_pos)
    trunc = good[:100] + good[-16:]                                   # metadata position points past the body
    with pytest.raises(ValueError):
        open_legacy_ivf_pq(trunc)
    with pytest.raises(ValueError, match="missing implementation"):
        parse_index_proto(_ld(1, b"x") + _ld(2, b"vector"))


def _four_bit(data):
    d = IvfPqData(**{**data.__dict__})
    d.codebook = data.codebook[:, :16, :].copy()
    return d


# ---- the library's native reader (csrc/api_file.inl) must agree with the Python one byte for byte -------------------
@pytest.mark.parametrize("key", sorted(FILES))
def test_native_reader_on_reference_written_files(key):
    from lance_b200.reader import open_legacy_ivf_pq_native
    want = open_legacy_ivf_pq(_path(*key))
    got = open_legacy_ivf_pq_native(_path(*key))
    assert (got.metric, got.dim, got.nlist, got.M) == (want.data.metric, want.data.dim, want.data.nlist, want.data.M)
    assert (got.name, got.column, got.dataset_version) == (want.meta.name, "vector", want.meta.dataset_version)
    for f in ("centroids", "codebook", "part_offsets", "codes", "row_ids"):
        assert np.array_equal(getattr(got, f), getattr(want.data, f)), f


@pytest.mark.parametrize("tensors", [True, False])
@pytest.mark.parametrize("metric", ["l2", "cosine", "dot"])
def test_native_reader_round_trip(metric, tensors):
    from lance_b200.reader import open_legacy_ivf_pq_native
    data = _synthetic(metric)
    got = open_legacy_ivf_pq_native(write_legacy(data, tensors=tensors))
    assert (got.metric, got.dim, got.nlist, got.M) == (metric, data.dim, data.nlist, data.M)
    for f in ("centroids", "codebook", "part_offsets", "codes", "row_ids"):
        assert np.array_equal(getattr(got, f), getattr(data, f)), f


def test_native_reader_errors(tmp_path):
    from lance_b200 import LanceB200Error
    from lance_b200.reader import open_legacy_ivf_pq_native
    data = _synthetic("l2")
    good = write_legacy(data)
    cases = [(good[-8:], "sufficient data"), (good[:-4] + b"XXXX", "magic"),
             (write_legacy(data, version=(0, 3)), "legacy IVF_PQ layout"),
             (write_legacy(_four_bit(data), num_bits=4), "8-bit"),
             (write_legacy(data, metric_code=3), "not supported"),
             (good[:-16] + struct.pack("<QHH", len(good), 0, 1) + b"LANC", "too small"),
             (good[:100] + good[-16:], None)]
    for blob, msg in cases:
        with pytest.raises(LanceB200Error, match=msg):
            open_legacy_ivf_pq_native(blob)
    with pytest.raises(LanceB200Error, match="cannot open"):
        open_legacy_ivf_pq_native(str(tmp_path / "missing.idx"))
    # a corrupted varint / truncated message inside the metadata must fail cleanly, never crash
    pos = struct.unpack("<Q", good[-16:-8])[0]
    rng = np.random.default_rng(0)
    for _ in range(200):
        b = bytearray(good)
        i = int(rng.integers(pos + 4, pos + 4 + 60))
        b[i] = int(rng.integers(0, 256))
        try:
            open_legacy_ivf_pq_native(bytes(b))
        except LanceB200Error:
            pass


def test_native_reader_survives_truncations_and_tail_edits():
    """Random truncations of a reference-written file (with and without a re-attached tail) and random edits of the
    16-byte tail: the native parser must either succeed or fail with an error code, never read out of bounds."""
    from lance_b200 import LanceB200Error
    from lance_b200.reader import open_legacy_ivf_pq_native
    good = open(_path("v0.21.0", "dc833a6e"), "rb").read()
    rng = np.random.default_rng(7)
    ok = 0
    for i in range(300):
        cut = int(rng.integers(0, len(good)))
        blob = good[:cut] + (good[-16:] if i % 2 else b"")
        if i % 3 == 0:
            t = bytearray(good[-16:])
            t[int(rng.integers(0, 12))] = int(rng.integers(0, 256))
            blob = good[:-16] + bytes(t)
        try:
            open_legacy_ivf_pq_native(blob)
            ok += 1
        except LanceB200Error:
            pass
    assert ok < 300


def test_native_reader_crafted_footer_and_sizes():
    """sizes taken from the file must be validated without wrapping additions and BEFORE they size an
    allocation.  Footer positions near 2^64 / at the tail, partition lengths and tensor shapes of 2^62: every one must
    come back as an error code (no out-of-bounds read, no std::bad_alloc crossing the C ABI)."""
    from lance_b200 import LanceB200Error
    from lance_b200.reader import open_legacy_ivf_pq_native
    data = _synthetic("l2")
    good = write_legacy(data)
    size = len(good)
    for pos in (2**64 - 1, 2**64 - 4, 2**64 - 20, size - 16, size - 19, size - 20, size, 2**63):
        blob = good[:-16] + struct.pack("<QHH", pos, 0, 1) + b"LANC"
        with pytest.raises(LanceB200Error):
            open_legacy_ivf_pq_native(blob)

    def with_ivf(lengths, offsets=None, cshape=None):
        offs = offsets if offsets is not None else [0] * len(lengths)
        ivf = _ld(2, b"".join(_vi(o) for o in offs)) + _ld(3, b"".join(_vi(n) for n in lengths))
        ct = _tensor(data.centroids)
        if cshape is not None:
            ct = _vf(1, 2) + _ld(2, b"".join(_vi(x) for x in cshape)) + _ld(3, b"")
        ivf += _ld(4, ct)
        pq = _vf(1, 8) + _vf(2, data.M) + _vf(3, data.dim) + _ld(5, _tensor(data.codebook.reshape(-1, data.dim // data.M)))
        vec = _vf(1, 1) + _vf(2, data.dim) + _ld(3, _ld(2, ivf)) + _ld(3, _ld(3, pq)) + _vf(4, 0)
        msg = _ld(1, b"i") + _ld(2, b"vector") + _vf(3, 7) + _vf(4, 0) + _ld(5, vec)
        body = b"\0" * 4096
        return body + struct.pack("<I", len(msg)) + msg + struct.pack("<QHH", len(body), 0, 1) + b"LANC"

    bad = [with_ivf([2**62] * 5), with_ivf([2**63, 2**63, 1, 1, 1]), with_ivf([100] * 5),   # sum exceeds the pages
           with_ivf([1] * 5, offsets=[2**64 - 8] * 5),
           with_ivf([1] * 5, cshape=[2**62, 4]), with_ivf([1] * 5, cshape=[2**32, 2**32]), with_ivf([1] * 5, cshape=[2**63, 2])]
    for blob in bad:
        with pytest.raises(LanceB200Error):
            open_legacy_ivf_pq_native(blob)
