"""`ME.utils` (reference: MinkowskiEngine/utils/__init__.py)."""
from .collation import SparseCollation, batch_sparse_collate, batched_coordinates, sparse_collate  # noqa: F401
from .init import kaiming_normal_  # noqa: F401
from .quantization import fnv_hash_vec, ravel_hash_vec, sparse_quantize, unique_coordinate_map  # noqa: F401
