"""v3 IVF_PQ index files (index.idx + auxiliary.idx, Lance-v2 containers): the Python reader and the
library's native reader must agree byte for byte on files written by tests/lance_v3_writer.py (a restatement of the
reference's writer -- PARITY UNPINNED: the reference tree holds no v3 fixture), and fail cleanly on malformed files."""
import os
import struct
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from lance_v3_writer import write_v3  # noqa: E402

from lance_b200.reader_v3 import open_v3_ivf_pq  # noqa: E402
from lance_b200.synth import IvfPqData  # noqa: E402


def _synthetic(metric="cosine", d=24, M=3, sizes=(7, 0, 33, 1, 16, 300), seed=3):
    rng = np.random.default_rng(seed)
    sizes = np.array(sizes)
    nlist = len(sizes)
    offs = np.zeros(nlist + 1, dtype=np.uint64)
    offs[1:] = np.cumsum(sizes)
    N = int(sizes.sum())
    return IvfPqData(metric, d, nlist, M, rng.normal(size=(nlist, d)).astype(np.float32),
                     rng.normal(size=(M, 256, d // M)).astype(np.float32), offs,
                     rng.integers(0, 256, size=N * M, dtype=np.uint8), (rng.permutation(1 << 20)[:N].astype(np.uint64) << np.uint64(21)) + np.uint64(5))


def _same(got, data):
    assert (got.metric, got.dim, got.nlist, got.M) == (data.metric, data.dim, data.nlist, data.M)
    for f in ("centroids", "codebook", "part_offsets", "codes", "row_ids"):
        assert np.array_equal(getattr(got, f), getattr(data, f)), f


@pytest.mark.parametrize("native", [False, True])
@pytest.mark.parametrize("transposed,rows_per_page,version", [(True, None, (0, 3)), (False, None, (2, 0)), (True, 5, (0, 3)),
                                                             (False, 7, (0, 3))])
@pytest.mark.parametrize("metric", ["l2", "cosine", "dot"])
def test_v3_round_trip(metric, transposed, rows_per_page, version, native, tmp_path):
    data = _synthetic(metric)
    idx, aux = write_v3(data, transposed=transposed, rows_per_page=rows_per_page, version=version)
    if native:
        from lance_b200.reader import open_v3_ivf_pq_native
        _same(open_v3_ivf_pq_native(idx, aux), data)
        (tmp_path / "index.idx").write_bytes(idx)
        (tmp_path / "auxiliary.idx").write_bytes(aux)
        _same(open_v3_ivf_pq_native(str(tmp_path / "index.idx"), str(tmp_path / "auxiliary.idx")), data)
    else:
        _same(open_v3_ivf_pq(idx, aux), data)


def test_v3_page_buffers_are_aligned_and_padded():
    data = _synthetic("l2")
    idx, aux = write_v3(data)
    assert idx[-4:] == b"LANC" and aux[-4:] == b"LANC"
    cm_start, cmo, gbo, ngb, ncol, major, minor = struct.unpack_from("<QQQIIHH", aux, len(aux) - 40)
    assert (ngb, ncol, major, minor) == (2, 2, 0, 3) and cm_start <= cmo <= gbo


def _errors():
    from lance_b200 import LanceB200Error
    return (ValueError, KeyError, IndexError, struct.error, LanceB200Error)


@pytest.mark.parametrize("native", [False, True])
def test_v3_malformed_files_fail_cleanly(native):
    from lance_b200.reader import open_v3_ivf_pq_native
    op = open_v3_ivf_pq_native if native else open_v3_ivf_pq
    data = _synthetic("l2")
    idx, aux = write_v3(data)
    errs = _errors()
    four = write_v3(data, num_bits=4)
    cases = [(idx[:-4] + b"XXXX", aux), (idx, aux[:-4] + b"XXXX"), (idx[:30], aux), (idx, aux[:39]),
             write_v3(data, version=(0, 1)), (aux, idx), four, (idx, write_v3(_synthetic("l2", sizes=(7, 0, 33)))[1])]
    for a, b in cases:
        with pytest.raises(errs):
            op(a, b)
    # crafted footers: offsets past the end / wrapping, table sizes that do not match the counts
    n = len(aux)
    for cm, cmo, gbo, ngb, ncol in [(2**64 - 1, 2**64 - 1, 2**64 - 1, 2, 2), (0, n, n, 2, 2), (0, 0, 0, 2**31, 2), (0, 0, 16, 1, 2**31),
                                    (n - 40, n - 40, n - 40, 0, 0)]:
        bad = aux[:-40] + struct.pack("<QQQIIHH", cm, cmo, gbo, ngb, ncol, 0, 3) + b"LANC"
        with pytest.raises(errs):
            op(idx, bad)
    # random corruption of the metadata region and truncations: an error or a (wrong but in-bounds) result, never a crash
    rng = np.random.default_rng(1)
    cm_start = struct.unpack_from("<Q", aux, n - 40)[0]
    gb0 = struct.unpack_from("<Q", aux, struct.unpack_from("<Q", aux, n - 24)[0])[0]
    for it in range(300):
        b = bytearray(aux)
        lo = min(cm_start, gb0)
        for _ in range(1 + it % 4):
            b[int(rng.integers(lo, n - 4))] = int(rng.integers(0, 256))
        if it % 5 == 0:
            b = b[:int(rng.integers(0, n))] + b[-40:]
        try:
            op(idx, bytes(b))
        except errs:
            pass
