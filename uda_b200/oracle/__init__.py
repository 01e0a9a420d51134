"""Pure-Python / NumPy oracles. Every CUDA kernel and every C++ runtime routine is
tested against a function in this package (SURVEY §4 "implications" item 1)."""
from .compare import get_compare_key, key_class_kind, SUPPORTED_KEY_CLASSES, UnsupportedKeyClass  # noqa: F401
from .merge import kway_merge, merged_stream, parse_stream  # noqa: F401
from .partition import range_partition, hash_partition, make_splitters  # noqa: F401
from .validate import teravalidate, record_checksum  # noqa: F401
