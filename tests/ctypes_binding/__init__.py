"""ctypes binding of libb200kv.so: test helper (the product binding is the compiled csrc/torch_binding.cpp)."""
from . import cache_ops, cuda_utils, ops  # noqa: F401
