"""lance_b200 -- B200-native IVF_PQ query path behind Lance's VectorIndex surface.

Python here is the thin host-side mirror of the reference interface used by tests and the
benchmark; the product is liblance_b200.so (CUDA, sm_100a) behind include/lance_b200.h.
"""
from .index import (DistanceType, IvfPqGroup, IvfPqIndex, IvfPqIndexF16, Prefilter, Query, VectorColumn, compute_distance, compute_distance_f16, flat_knn,  # noqa: F401
                    flat_knn_f16, merge_topk, normalize, rerank)
from ._lib import LanceB200Error, SO_PATH  # noqa: F401

# Callers either side of the index: `lance_b200.reader` (legacy index.idx -> device index, Python and native
# parsers), `lance_b200.scanner.VectorDataset` (Scanner::nearest plan), `lance_b200.shard` (multi-GPU).
