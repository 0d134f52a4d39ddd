"""ctypes loader for liblance_b200.so (the C ABI declared in include/lance_b200.h).

There is no fallback of any kind: if the shared library is missing the import fails with an
explicit build hint, and every compute call fails loudly when no CUDA device is usable.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.environ.get("LB200_SO") or os.path.join(_HERE, "liblance_b200.so")  # LB200_SO: A/B of builds

EXPORTS = [
    "lb200_index_create", "lb200_index_destroy", "lb200_find_partitions", "lb200_search_partitions",
    "lb200_search", "lb200_search_device", "lb200_find_partitions_device", "lb200_search_probed_device", "lb200_flat_distance", "lb200_flat_topk",
    "lb200_flat_topk_device", "lb200_merge_topk_device", "lb200_set_profiling", "lb200_get_stats",
    "lb200_set_scan_variant", "lb200_last_error", "lb200_version",
    "lb200_peer_thresholds_export", "lb200_peer_thresholds_attach", "lb200_peer_thresholds_detach",
    "lb200_merge_topk", "lb200_rerank",
    "lb200_index_file_open", "lb200_index_file_open_bytes", "lb200_index_file_get_info", "lb200_index_file_close",
    "lb200_index_open", "lb200_flat_distance_f16", "lb200_flat_topk_f16", "lb200_normalize",
    "lb200_shard_export", "lb200_shard_attach", "lb200_shard_detach", "lb200_shard_search", "lb200_shard_search_device",
    "lb200_shard_check", "lb200_group_create", "lb200_group_search", "lb200_group_destroy",
    "lb200_index_file_open_v3", "lb200_index_file_open_v3_bytes", "lb200_index_open_v3",
    "lb200_kmeans_train", "lb200_kmeans_train_device", "lb200_compute_partitions", "lb200_compute_partitions_device",
    "lb200_residuals_device", "lb200_pq_train_device", "lb200_pq_encode", "lb200_pq_encode_device",
    "lb200_column_create", "lb200_column_destroy", "lb200_column_flat_topk", "lb200_column_flat_topk_device",
    "lb200_column_last_call",
    "lb200_flat_topk_device_noalloc", "lb200_noalloc_scratch_bytes",
    "lb200_prefilter_create", "lb200_prefilter_destroy", "lb200_search_partitions_filtered",
    "lb200_index_create_f16", "lb200_search_f16", "lb200_find_partitions_f16", "lb200_search_partitions_f16",
]


class Stats(C.Structure):
    _fields_ = [
        ("scan_ms", C.c_double), ("coarse_ms", C.c_double), ("total_ms", C.c_double),
        ("scan_code_bytes", C.c_uint64), ("pairs", C.c_uint64), ("kernel_launches", C.c_uint32),
        ("scan_variant", C.c_uint32), ("scan_ql", C.c_uint32), ("scan_items", C.c_uint32),
    ]


class LanceB200Error(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"lance_b200 error {code}: {msg}")
        self.code = code


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(SO_PATH):
            raise ImportError(
                f"{SO_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "or `make -C lance_b200/csrc` (nvcc, sm_100a). lance_b200 has no CPU fallback.")
        _lib = C.CDLL(SO_PATH)
        _lib.lb200_last_error.restype = C.c_char_p
        _lib.lb200_version.restype = C.c_char_p
        _lib.lb200_index_destroy.restype = None
        _lib.lb200_prefilter_destroy.restype = None
        _lib.lb200_column_destroy.restype = None
        _lib.lb200_noalloc_scratch_bytes.restype = C.c_uint64
        _lib.lb200_group_destroy.restype = None
        for name in EXPORTS:
            if os.environ.get("LB200_SO") and not hasattr(_lib, name):
                continue  # an older build under A/B test may predate a symbol
            getattr(_lib, name)  # raises AttributeError if a declared symbol is missing
    return _lib


def check(rc):
    if rc != 0:
        raise LanceB200Error(rc, lib().lb200_last_error().decode())
