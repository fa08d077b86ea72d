"""`vllm._C`-shaped module: submodules `ops`, `cache_ops`, `cuda_utils` with the reference's
positional signatures (csrc/pybind.cpp:6-129), implemented over libb200kv.so."""
from .. import _lib

_lib.load()  # fail loudly at import time if the CUDA library is missing

from . import cache_ops, cuda_utils, ops  # noqa: E402

__all__ = ["ops", "cache_ops", "cuda_utils"]
