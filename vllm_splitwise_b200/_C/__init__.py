"""`vllm._C`-shaped module: submodules `ops`, `cache_ops`, `cuda_utils` with the reference's positional
signatures (csrc/pybind.cpp:6-129), implemented over libb200kv.so.

The binding is lib/_b200_torch.so, the compiled pybind11 front (csrc/torch_binding.cpp) over the C ABI of
include/b200_kv.h: it only marshals tensors into pointers, sizes and strides.  The kernels are the CUDA ones in
libb200kv.so; a missing library or a missing front raises at import time -- there is no fallback.
(tests/ctypes_binding/ holds a second binding of the same ABI through ctypes, as a test helper.)
"""
import importlib.util
import os
import sys

from .. import _lib

_lib.load()  # fail loudly at import time if the CUDA library is missing


def _load_compiled():
    path = os.path.join(os.path.dirname(_lib.LIB_PATH), "_b200_torch.so")
    if not os.path.exists(path):
        raise ImportError(f"{path} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(or bash vllm_splitwise_b200/csrc/build_torch_binding.sh). There is no fallback binding.")
    import torch  # noqa: F401  (libtorch must be loaded before the extension)
    spec = importlib.util.spec_from_file_location("_b200_torch", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


_mod = _load_compiled()
ops, cache_ops, cuda_utils = _mod.ops, _mod.cache_ops, _mod.cuda_utils
BINDING = "torch"
# `import vllm_splitwise_b200._C.ops` style imports resolve to the binding too
for _name, _m in (("ops", ops), ("cache_ops", cache_ops), ("cuda_utils", cuda_utils)):
    sys.modules[__name__ + "." + _name] = _m

__all__ = ["ops", "cache_ops", "cuda_utils", "BINDING"]
