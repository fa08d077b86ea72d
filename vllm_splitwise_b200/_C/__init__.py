"""`vllm._C`-shaped module: submodules `ops`, `cache_ops`, `cuda_utils` with the reference's
positional signatures (csrc/pybind.cpp:6-129), implemented over libb200kv.so.

Two bindings of the same C ABI (include/b200_kv.h) exist; both only marshal arguments:
  * lib/_b200_torch.so -- compiled pybind11 front (csrc/torch_binding.cpp), ~1 us per call; preferred.
  * _ctypes_*.py       -- the same binding through ctypes, used when the compiled front has not been
                          built (or B200_KV_BINDING=ctypes).
Either way the kernels are the CUDA ones in libb200kv.so; a missing library raises at import time.
"""
import importlib.util
import os
import sys

from .. import _lib

_lib.load()  # fail loudly at import time if the CUDA library is missing

BINDING = None


def _load_compiled():
    path = os.path.join(os.path.dirname(_lib.LIB_PATH), "_b200_torch.so")
    if not os.path.exists(path):
        return None
    import torch  # noqa: F401  (libtorch must be loaded before the extension)
    spec = importlib.util.spec_from_file_location("_b200_torch", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


_want = os.environ.get("B200_KV_BINDING", "auto")
_mod = _load_compiled() if _want in ("auto", "torch") else None
if _mod is None and _want == "torch":
    raise ImportError("B200_KV_BINDING=torch but lib/_b200_torch.so is not built "
                      "(bash vllm_splitwise_b200/csrc/build_torch_binding.sh)")
if _mod is not None:
    ops, cache_ops, cuda_utils = _mod.ops, _mod.cache_ops, _mod.cuda_utils
    BINDING = "torch"
else:
    from . import _ctypes_cache_ops as cache_ops, _ctypes_cuda_utils as cuda_utils, _ctypes_ops as ops  # noqa: E402
    BINDING = "ctypes"
# `import vllm_splitwise_b200._C.ops` style imports resolve to the selected binding too
for _name, _m in (("ops", ops), ("cache_ops", cache_ops), ("cuda_utils", cuda_utils)):
    sys.modules[__name__ + "." + _name] = _m

__all__ = ["ops", "cache_ops", "cuda_utils", "BINDING"]
