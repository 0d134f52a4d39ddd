"""lb200_flat_topk_device_noalloc (no counterpart in the reference): the
flat kNN top-k with its scratch in a fixed `static __device__` array, for callers that cannot call cudaMalloc.  Called
through the C ABI twice in a row on different inputs and compared with the existing lb200_flat_topk_device -- the bar of
the other GPU tests: identical indices and bit-identical distances.  Collected last on
purpose."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _run(L, fn, noalloc, metric, q, v, valid, k, stream):
    import torch
    dev = torch.device("cuda", 0)
    Q, d = q.shape
    n = v.shape[0]
    d_q, d_v = torch.from_numpy(q).to(dev), torch.from_numpy(v).to(dev)
    d_m = torch.from_numpy(valid).to(dev) if valid is not None else None
    od = torch.full((Q, k), -1.0, dtype=torch.float32, device=dev)
    oi = torch.full((Q, k), -7, dtype=torch.int64, device=dev)
    oc = torch.full((Q,), -1, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    rc = fn(C.c_int32(0), C.c_int32(metric), C.c_uint32(d), C.c_void_p(d_q.data_ptr()), C.c_uint32(Q),
            C.c_void_p(d_v.data_ptr()), C.c_void_p(d_m.data_ptr()) if d_m is not None else None, C.c_uint64(n), C.c_uint32(k),
            C.c_void_p(od.data_ptr()), C.c_void_p(oi.data_ptr()), C.c_void_p(oc.data_ptr()), C.c_void_p(stream.cuda_stream))
    assert rc == 0, L.lb200_last_error().decode()
    stream.synchronize()
    return od.cpu().numpy(), oi.cpu().numpy(), oc.cpu().numpy()


def test_noalloc_flat_topk_equals_flat_topk_device_twice_in_a_row():
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    from lance_b200 import _lib
    L = _lib.lib()
    assert L.lb200_noalloc_scratch_bytes() == 32 << 20
    stream = torch.cuda.Stream(device=torch.device("cuda", 0))
    rng = np.random.default_rng(17)
    # call 1: several select segments + merge, validity bitmap, L2;  call 2 (smaller, other metric, one segment) must
    # not see anything of call 1 in the shared scratch;  call 3: the first shape again
    cases = [(0, 9, 128, 150_001, 10, True), (2, 5, 70, 3_001, 100, False), (1, 16, 64, 70_000, 1, True),
             (0, 9, 128, 150_001, 10, True)]
    for metric, Q, d, n, k, with_valid in cases:
        base = rng.normal(size=(997, d)).astype(np.float32)
        v = base[rng.integers(0, 997, n)]                      # duplicated rows: ties across segment borders
        q = base[:Q] + 0.01 * rng.normal(size=(Q, d)).astype(np.float32)
        valid = np.packbits(rng.random(n) < 0.95, bitorder="little") if with_valid else None
        want = _run(L, L.lb200_flat_topk_device, False, metric, q, v, valid, k, stream)
        got = _run(L, L.lb200_flat_topk_device_noalloc, True, metric, q, v, valid, k, stream)
        assert got[2].tolist() == want[2].tolist()
        assert got[1].tolist() == want[1].tolist()
        assert got[0].view(np.uint32).tolist() == want[0].view(np.uint32).tolist()
        assert np.all(got[2] == min(k, n)) or with_valid


def test_noalloc_rejects_sizes_beyond_the_fixed_scratch():
    import torch
    from lance_b200 import _lib
    L = _lib.lib()
    dev = torch.device("cuda", 0)
    Q, d, n, k = 64, 16, 200_000, 10                            # 64 * 200k * 4 = 51 MB > 32 MiB
    d_q = torch.zeros((Q, d), device=dev)
    d_v = torch.zeros((n, d), device=dev)
    od = torch.zeros((Q, k), device=dev)
    oi = torch.zeros((Q, k), dtype=torch.int64, device=dev)
    oc = torch.zeros((Q,), dtype=torch.int32, device=dev)
    rc = L.lb200_flat_topk_device_noalloc(C.c_int32(0), C.c_int32(0), C.c_uint32(d), C.c_void_p(d_q.data_ptr()), C.c_uint32(Q),
                                          C.c_void_p(d_v.data_ptr()), None, C.c_uint64(n), C.c_uint32(k), C.c_void_p(od.data_ptr()),
                                          C.c_void_p(oi.data_ptr()), C.c_void_p(oc.data_ptr()), None)
    assert rc == -1
    msg = L.lb200_last_error().decode()
    assert "scratch" in msg and str(32 << 20) in msg
    # n = 0: empty results, no scratch
    rc = L.lb200_flat_topk_device_noalloc(C.c_int32(0), C.c_int32(0), C.c_uint32(d), C.c_void_p(d_q.data_ptr()), C.c_uint32(Q),
                                          None, None, C.c_uint64(0), C.c_uint32(k), C.c_void_p(od.data_ptr()),
                                          C.c_void_p(oi.data_ptr()), C.c_void_p(oc.data_ptr()), None)
    torch.cuda.synchronize()
    assert rc == 0 and oc.cpu().tolist() == [0] * Q
