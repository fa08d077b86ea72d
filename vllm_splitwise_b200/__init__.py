"""Importable alias of the `vllm-splitwise_b200/` package directory.

The product lives in `vllm-splitwise_b200/` (a hyphen is not a legal Python identifier); this
module gives it the import name `vllm_splitwise_b200` by pointing `__path__` at that directory.
"""
import os as _os

_pkg_dir = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))),
                         "vllm-splitwise_b200")
__path__ = [_pkg_dir]
__file__ = _os.path.join(_pkg_dir, "__init__.py")
with open(__file__) as _f:
    exec(compile(_f.read(), __file__, "exec"))
