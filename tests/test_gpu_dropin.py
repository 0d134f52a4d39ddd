"""The drop-in call pattern: the exec nodes call find_partitions once and search_in_partition
once per probed partition with several calls in flight (rust/lance/src/io/exec/knn.rs:349-393,606-634).  These tests
pin the small-call fast path (work list built on the host), the device-resident prefilter and concurrent use of one
handle against the batched path / the oracle, bit for bit."""
import os
import threading

import numpy as np
import pytest
import torch

from lance_b200.synth import random_index, queries_near_centroids

pytestmark = pytest.mark.gpu


def _bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


@pytest.fixture(scope="module")
def setup():
    assert torch.cuda.is_available()
    from lance_b200 import IvfPqIndex
    data = random_index(300_000, 128, 256, 16, "l2", seed=5)
    q = queries_near_centroids(data, 64, seed=6)
    ix = IvfPqIndex(data)
    yield data, q, ix
    ix.close()


def test_small_calls_equal_batched_call_and_oracle(oracle, setup):
    data, q, ix = setup
    probes = ix.find_partitions(q, 8)
    want_d, want_i, want_c = ix.search_in_partition(probes, q, 10)          # one batched call (device work list)
    o = oracle.OracleIndex(data)
    for qi in range(0, 64, 7):
        one = ix.find_partitions(q[qi], 8)                                   # Q = 1
        assert one.tolist() == probes[qi:qi + 1].tolist()
        for j in range(8):
            d, i, c = ix.search_in_partition(probes[qi:qi + 1, j:j + 1], q[qi], 10)   # Q = 1, P = 1 (host work list)
            n = int(c[0, 0])
            assert n == want_c[qi, j]
            assert i[0, 0, :n].tolist() == want_i[qi, j, :n].tolist()
            assert _bits(d[0, 0, :n]).tolist() == _bits(want_d[qi, j, :n]).tolist()
            rd, ri = o.search_in_partition(int(probes[qi, j]), q[qi], 10)
            assert i[0, 0, :n].tolist() == ri.tolist() and _bits(d[0, 0, :n]).tolist() == _bits(rd).tolist()


def test_host_and_device_work_lists_agree(setup):
    data, q, ix = setup
    probes = ix.find_partitions(q[:16], 5)
    a = ix.search_in_partition(probes, q[:16], 10)          # 80 pairs: host-built work list
    os.environ["LB200_NO_HOST_WORKLIST"] = "1"
    try:
        b = ix.search_in_partition(probes, q[:16], 10)      # work-list kernels
    finally:
        del os.environ["LB200_NO_HOST_WORKLIST"]
    assert a[2].tolist() == b[2].tolist() and a[1].tolist() == b[1].tolist()
    assert _bits(a[0]).tolist() == _bits(b[0]).tolist()


def test_prefilter_object_equals_host_masks(oracle, setup):
    data, q, ix = setup
    rng = np.random.default_rng(3)
    probes = ix.find_partitions(q[:8], 6)
    sizes = np.diff(data.part_offsets.astype(np.int64))
    masks = {int(p): rng.random(int(sizes[p])) < 0.25 for p in np.unique(probes)[::2]}
    want = ix.search_in_partition(probes, q[:8], 10, prefilter=masks)
    pf = ix.prefilter(masks)
    got = ix.search_in_partition(probes, q[:8], 10, prefilter=pf)
    assert got[2].tolist() == want[2].tolist() and got[1].tolist() == want[1].tolist()
    assert _bits(got[0]).tolist() == _bits(want[0]).tolist()
    o = oracle.OracleIndex(data)
    for qi in range(8):                                       # single-pair calls share the same device masks
        for j in range(6):
            p = int(probes[qi, j])
            d, i, c = ix.search_in_partition(probes[qi:qi + 1, j:j + 1], q[qi], 10, prefilter=pf)
            bm = None
            if p in masks:
                bm = np.zeros((int(sizes[p]) + 63) // 64 + 1, dtype=np.uint64)
                packed = np.packbits(masks[p], bitorder="little")
                bm.view(np.uint8)[:packed.size] = packed
            rd, ri = o.search_in_partition(p, q[qi], 10, bitmask=bm)
            n = len(rd)
            assert c[0, 0] == n and i[0, 0, :n].tolist() == ri.tolist()
            assert _bits(d[0, 0, :n]).tolist() == _bits(rd).tolist()
    pf.close()


def test_concurrent_calls_on_one_handle(setup):
    """16 host threads issue single-pair searches, batched searches and find_partitions on ONE handle at once;
    every result must equal the serial one (per-call contexts: no shared scratch, no serialising mutex)."""
    from lance_b200 import Query
    data, q, ix = setup
    probes = ix.find_partitions(q, 8)
    want = ix.search_in_partition(probes, q, 10)
    want_s = ix.search(Query(key=q, k=10, nprobes=8))
    errors = []

    def worker(t):
        try:
            rng = np.random.default_rng(t)
            for it in range(60):
                qi, j = int(rng.integers(0, 64)), int(rng.integers(0, 8))
                kind = it % 4
                if kind < 2:
                    d, i, c = ix.search_in_partition(probes[qi:qi + 1, j:j + 1], q[qi], 10)
                    n = int(c[0, 0])
                    assert n == want[2][qi, j] and i[0, 0, :n].tolist() == want[1][qi, j, :n].tolist()
                    assert _bits(d[0, 0, :n]).tolist() == _bits(want[0][qi, j, :n]).tolist()
                elif kind == 2:
                    assert ix.find_partitions(q[qi], 8).tolist() == probes[qi:qi + 1].tolist()
                else:
                    lo = int(rng.integers(0, 48))
                    d, i, c = ix.search(Query(key=q[lo:lo + 16], k=10, nprobes=8))
                    assert i.tolist() == want_s[1][lo:lo + 16].tolist()
                    assert _bits(d).tolist() == _bits(want_s[0][lo:lo + 16]).tolist()
        except Exception as e:  # noqa: BLE001
            errors.append(repr(e))

    threads = [threading.Thread(target=worker, args=(t,)) for t in range(16)]
    for th in threads:
        th.start()
    for th in threads:
        th.join()
    assert not errors, errors[:3]
