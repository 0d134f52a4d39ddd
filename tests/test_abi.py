"""CPU-only checks of the drop-in boundary: the C-ABI library loads and exports every symbol
that include/lance_b200.h declares; argument validation that needs no GPU behaves like the
reference's; without a GPU compute calls fail loudly (no CPU fallback)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    h = open(os.path.join(ROOT, "include", "lance_b200.h")).read()
    h = re.sub(r"/\*.*?\*/", "", h, flags=re.S)
    return sorted(set(re.findall(r"\b(lb200_[a-z0-9_]+)\s*\(", h)))


def test_header_symbols_exported():
    from lance_b200 import _lib
    L = _lib.lib()
    names = _declared()
    assert len(names) >= 15
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/lance_b200.h but not exported"
    assert sorted(_lib.EXPORTS) == names
    assert b"sm_100a" in L.lb200_version()


def test_no_oracle_in_product():
    """The product path must never touch oracle/ (prompt 3): no import, no dlopen, no link."""
    pkg = os.path.join(ROOT, "lance_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".inl", ".h", "Makefile")):
                src = open(os.path.join(dp, f), errors="ignore").read()
                assert not re.search(r"import\s+oracle|from\s+oracle|liblance_oracle|oracle/|lo_[a-z_]+\(", src), \
                    f"{f} references the oracle"


def test_argument_validation_without_gpu():
    from lance_b200 import _lib
    L = _lib.lib()
    h = C.c_void_p()
    cent = np.zeros((2, 4), np.float32)
    cb = np.zeros((2, 256, 2), np.float32)
    offs = np.zeros(3, np.uint64)
    args = lambda bits, M=2, metric=0: (C.c_int32(0), C.c_int32(metric), C.c_uint32(4), C.c_uint32(2),
                                        cent.ctypes.data_as(C.c_void_p), C.c_uint32(M), C.c_uint32(bits),
                                        cb.ctypes.data_as(C.c_void_p), offs.ctypes.data_as(C.c_void_p), None, None,
                                        None, C.byref(h))
    assert L.lb200_index_create(*args(5)) == -4  # only 8-bit and 4-bit PQ exist (pq.rs:103-111)
    assert b"8-bit and 4-bit" in L.lb200_last_error()
    assert L.lb200_index_create(*args(4, M=1)) == -1  # 4-bit: num_sub_vectors must be even (pq.rs:131-138)
    assert b"divisible by 2" in L.lb200_last_error()
    assert L.lb200_index_create(*args(8, M=3)) == -1  # dim % M != 0
    assert L.lb200_index_create(*args(8, metric=7)) == -1
    assert L.lb200_set_scan_variant(None, 0) == -1


def test_fails_loudly_without_cuda():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from lance_b200 import _lib
    L = _lib.lib()
    h = C.c_void_p()
    cent = np.zeros((2, 4), np.float32)
    cb = np.zeros((2, 256, 2), np.float32)
    offs = np.zeros(3, np.uint64)
    rc = L.lb200_index_create(C.c_int32(0), C.c_int32(0), C.c_uint32(4), C.c_uint32(2),
                              cent.ctypes.data_as(C.c_void_p), C.c_uint32(2), C.c_uint32(8),
                              cb.ctypes.data_as(C.c_void_p), offs.ctypes.data_as(C.c_void_p), None, None, None,
                              C.byref(h))
    assert rc == -2 and b"no CPU fallback" in L.lb200_last_error()
    with pytest.raises(_lib.LanceB200Error):
        from lance_b200 import flat_knn
        flat_knn(np.zeros(4, np.float32), "l2", np.zeros((3, 4), np.float32), 1)


def test_host_mirror_validation():
    from lance_b200 import DistanceType
    with pytest.raises(ValueError):
        DistanceType.code("manhattan")
    assert DistanceType.code("L2") == 0 and DistanceType.code("cosine") == 1 and DistanceType.code("dot") == 2


def test_header_is_plain_c_and_a_c_host_links(tmp_path):
    """include/lance_b200.h must be consumable from C (the reference-side binding is an `extern "C"` block in a
    Rust -sys crate): compile examples/open_legacy_index.c with gcc -std=c11, link it against the built library and
    run its host-only part on a reference-written index file."""
    import subprocess
    from conftest import lance_files_dir
    exe = str(tmp_path / "open_legacy_index")
    libdir = os.path.join(ROOT, "lance_b200")
    subprocess.check_call(["gcc", "-std=c11", "-Wall", "-Wextra", "-Werror", "-I", os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "examples", "open_legacy_index.c"), "-L", libdir, "-llance_b200",
                           "-Wl,-rpath," + libdir, "-o", exe])
    out = subprocess.run([exe, os.path.join(lance_files_dir(), "v0.8.14_f358a219_index.idx")], capture_output=True,
                         text=True, timeout=120)
    assert out.returncode == 0, out.stderr
    assert "sm_100a" in out.stdout
    assert "name=vector_idx column=vector metric=0 dim=128 nlist=4 M=16 bits=8 rows=2000" in out.stdout
