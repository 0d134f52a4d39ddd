"""Host-side restatements of two device layouts, checked for the properties the kernels rely on (no GPU needed).

* the lane-order codebook image of the in-kernel table build (csrc/scan3.cuh `cb_image_f4`): a permutation of the
  codebook's float4s in which the 32 lanes of one load instruction of the build read 512 contiguous bytes;
* the runs of the work-list prefix kernels (csrc/worklist.cuh `offsets_kernel`): every partition is owned by exactly one
  of the 1024 threads, in order, for any partition count.
"""
import numpy as np
import pytest


def cb_image_f4(ch, code, slot, i, dsub4):
    return ((ch * 128 + (code >> 1)) * dsub4 + i) * 32 + (code & 1) * 16 + slot


@pytest.mark.parametrize("M,dsub", [(16, 8), (48, 16), (96, 16), (32, 8)])
def test_codebook_image_is_a_permutation(M, dsub):
    dsub4 = dsub // 4
    m, code, i = np.meshgrid(np.arange(M), np.arange(256), np.arange(dsub4), indexing="ij")
    off = cb_image_f4(m >> 4, code, m & 15, i, dsub4).ravel()
    assert off.min() == 0 and off.max() == M * 256 * dsub4 - 1 and len(np.unique(off)) == off.size


@pytest.mark.parametrize("NT,dsub", [(256, 8), (256, 16), (512, 8), (512, 16)])
def test_codebook_image_loads_are_contiguous_per_warp(NT, dsub):
    """Build thread tid: slot = tid % 16, code = tid / 16 + step * NT / 16 (scan3.cuh A4).  For every step, chunk and
    float4 index the 32 lanes of a warp must read 32 consecutive float4s, starting at a 512-byte boundary."""
    dsub4 = dsub // 4
    tid = np.arange(NT)
    for ch in (0, 2):
        for step in range(256 // (NT // 16)):
            code = (tid >> 4) + step * (NT // 16)
            for i in range(dsub4):
                off = cb_image_f4(ch, code, tid & 15, i, dsub4).reshape(NT // 32, 32)
                assert (np.diff(off, axis=1) == 1).all() and (off[:, 0] % 32 == 0).all()
                # and the kernel's pointer arithmetic reaches the same place: base(tid) + step * CBSTEP + i * 32
                base = cb_image_f4(ch, tid >> 4, tid & 15, 0, dsub4)
                assert (base + step * (NT // 32) * dsub4 * 32 + i * 32 == off.ravel()).all()


@pytest.mark.parametrize("nlist", [1, 31, 1023, 1024, 1025, 1500, 2049, 4096, 65536])
def test_prefix_kernel_runs_cover_every_partition_once(nlist):
    per = (nlist + 1023) // 1024
    t = np.arange(1024)
    p0 = np.minimum(nlist, t * per)
    p1 = np.minimum(nlist, p0 + per)
    assert p0[0] == 0 and p1[-1] == nlist and (p1[:-1] == p0[1:]).all() and ((p1 - p0) <= per).all()
