"""lance_b200 -- B200-native IVF_PQ query path behind Lance's VectorIndex surface.

Python here is the thin host-side mirror of the reference interface used by tests and the
benchmark; the product is liblance_b200.so (CUDA, sm_100a) behind include/lance_b200.h.
"""
from .index import DistanceType, IvfPqIndex, Query, flat_knn, compute_distance  # noqa: F401
from ._lib import LanceB200Error, SO_PATH  # noqa: F401
