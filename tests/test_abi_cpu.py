"""CPU: the C-ABI library loads without a GPU and exports exactly what include/b200_kv.h declares;
argument validation that needs no device works; the product package never touches the oracle."""
import ctypes
import os
import re
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "b200_kv.h")


def _declared():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"B200_API\s+[\w\s\*]+?\b(b200_\w+)\s*\(", src)))


def test_header_symbols_exported_and_bound():
    from vllm_splitwise_b200 import _lib
    names = _declared()
    assert len(names) >= 19
    lib = _lib.load()
    for n in names:
        assert hasattr(lib, n), f"{n} declared in b200_kv.h but not exported by libb200kv.so"
    assert sorted(_lib.SIGNATURES) == names, "ctypes table and header disagree"
    out = subprocess.run(["nm", "-D", "--defined-only", _lib.LIB_PATH], capture_output=True, text=True, check=True)
    exported = sorted(l.split()[-1] for l in out.stdout.splitlines() if " T " in l)
    assert [e for e in exported if e.startswith("b200_")] == names
    # nothing else leaks out of the library (hidden visibility)
    assert all(e.startswith("b200_") or e.startswith("_") for e in exported), exported


def test_struct_layout_matches_header():
    """sizeof/offsetof of b200_paged_attn_args_t as compiled by gcc == the ctypes mirror."""
    from vllm_splitwise_b200 import _lib
    fields = [f[0] for f in _lib.PagedAttnArgs._fields_]
    prog = '#include <stdio.h>\n#include <stddef.h>\n#include "b200_kv.h"\nint main(){printf("%zu", sizeof(b200_paged_attn_args_t));' + \
        "".join(f'printf(" %zu", offsetof(b200_paged_attn_args_t, {f}));' for f in fields) + "return 0;}"
    exe = "/tmp/_b200_layout"
    subprocess.run(["gcc", "-x", "c", "-", "-I", os.path.join(ROOT, "include"), "-o", exe], input=prog, text=True, check=True)
    got = list(map(int, subprocess.run([exe], capture_output=True, text=True, check=True).stdout.split()))
    assert got[0] == ctypes.sizeof(_lib.PagedAttnArgs)
    assert got[1:] == [getattr(_lib.PagedAttnArgs, f).offset for f in fields]
    # b200_decode_step_args_t too
    fields = [f[0] for f in _lib.DecodeStepArgs._fields_]
    prog = '#include <stdio.h>\n#include <stddef.h>\n#include "b200_kv.h"\nint main(){printf("%zu", sizeof(b200_decode_step_args_t));' + \
        "".join(f'printf(" %zu", offsetof(b200_decode_step_args_t, {f}));' for f in fields) + "return 0;}"
    subprocess.run(["gcc", "-x", "c", "-", "-I", os.path.join(ROOT, "include"), "-o", exe], input=prog, text=True, check=True)
    got = list(map(int, subprocess.run([exe], capture_output=True, text=True, check=True).stdout.split()))
    assert got[0] == ctypes.sizeof(_lib.DecodeStepArgs)
    assert got[1:] == [getattr(_lib.DecodeStepArgs, f).offset for f in fields]


def test_status_strings_and_validation_without_gpu():
    from vllm_splitwise_b200 import _lib
    lib = _lib.load()
    assert lib.b200_kv_abi_version() == 2
    assert lib.b200_kv_strerror(_lib.ERR_HEAD_SIZE) == b"Unsupported head size"
    assert lib.b200_kv_strerror(_lib.ERR_BLOCK_SIZE) == b"Unsupported block size"
    assert lib.b200_kv_strerror(_lib.ERR_DTYPE) == b"Unsupported data type"
    a = _lib.PagedAttnArgs()
    assert lib.b200_paged_attention_v1(None, None) == _lib.ERR_ARG
    a.num_seqs = 0
    assert lib.b200_paged_attention_v1(ctypes.byref(a), None) == _lib.B200_OK  # empty batch: no launch
    a.num_seqs = 2
    assert lib.b200_paged_attention_v1(ctypes.byref(a), None) == _lib.ERR_ARG  # null pointers
    for f in ("out", "query", "key_cache", "value_cache", "block_tables", "context_lens"):
        setattr(a, f, 4096)
    a.num_heads, a.num_kv_heads, a.head_size, a.block_size = 8, 3, 128, 16
    assert lib.b200_paged_attention_v1(ctypes.byref(a), None) == _lib.ERR_ARG  # 8 % 3
    a.num_kv_heads, a.head_size = 2, 72
    assert lib.b200_paged_attention_v1(ctypes.byref(a), None) == _lib.ERR_HEAD_SIZE
    a.head_size, a.block_size = 128, 4
    assert lib.b200_paged_attention_v1(ctypes.byref(a), None) == _lib.ERR_BLOCK_SIZE
    a.block_size, a.query_dtype, a.cache_dtype = 16, _lib.U8_E5M2, _lib.U8_E5M2
    assert lib.b200_paged_attention_v1(ctypes.byref(a), None) == _lib.ERR_DTYPE
    a.query_dtype, a.cache_dtype = _lib.F16, _lib.BF16
    assert lib.b200_paged_attention_v1(ctypes.byref(a), None) == _lib.ERR_DTYPE
    a.cache_dtype, a.kv_block_stride, a.kv_head_stride = _lib.F16, 2048 * 2 + 1, 2048
    assert lib.b200_paged_attention_v1(ctypes.byref(a), None) == _lib.ERR_ALIGN
    assert lib.b200_reshape_and_cache(None, None, None, None, None, 0, 1, 1, 1, 1, 0, 0, 0, 0, 0, None) == _lib.B200_OK
    assert lib.b200_reshape_and_cache(None, None, None, None, None, 3, 1, 1, 1, 1, 0, 0, 0, 0, 0, None) == _lib.ERR_ARG
    assert lib.b200_copy_blocks(None, None, 0, None, 0, 0, None) == _lib.B200_OK
    assert lib.b200_swap_blocks(4096, 8192, 64, (ctypes.c_int64 * 2)(0, 1), 1, _lib.SWAP_D2D, 0, 1, None) == _lib.ERR_DEVICE
    assert lib.b200_kv_put_blocks(4096, 4096, 1, (ctypes.c_int32 * 2)(0, 1), 1, 24, 0, None) == _lib.ERR_ALIGN
    with pytest.raises(RuntimeError, match="Unsupported head size: 72"):
        _lib.check(_lib.ERR_HEAD_SIZE, "paged_attention_v1", 72)


def _bindings():
    """(name, ops, cache_ops) for every binding of the C ABI that is built: compiled pybind front, ctypes."""
    from vllm_splitwise_b200 import _C
    from ctypes_binding import cache_ops as _ctypes_cache_ops, ops as _ctypes_ops
    out = [("ctypes", _ctypes_ops, _ctypes_cache_ops)]
    if _C.BINDING == "torch":
        out.append(("torch", _C.ops, _C.cache_ops))
    return out


def test_compiled_binding_is_built_and_selected():
    """build() compiles lib/_b200_torch.so; `_C` must pick it (tests/ctypes_binding is a test helper)."""
    from vllm_splitwise_b200 import _C
    assert os.path.exists(os.path.join(ROOT, "vllm_splitwise_b200", "lib", "_b200_torch.so"))
    assert _C.BINDING == "torch"
    for sub, names in (("ops", ["paged_attention_v1", "paged_attention_v2"]),
                       ("cache_ops", ["reshape_and_cache", "copy_blocks", "swap_blocks", "convert_fp8_e5m2"]),
                       ("cuda_utils", ["get_device_attribute", "get_max_shared_memory_per_block_device_attribute"])):
        for n in names:
            assert callable(getattr(getattr(_C, sub), n)), (sub, n)


@pytest.mark.parametrize("which", ["ctypes", "torch"])
def test_ops_refuse_cpu_tensors(which):
    import torch
    found = {n: (o, c) for n, o, c in _bindings()}
    if which not in found:
        pytest.skip(f"{which} binding not built")
    ops, cache_ops = found[which]
    q = torch.zeros(1, 2, 64)
    with pytest.raises(RuntimeError, match="no CPU path"):
        ops.paged_attention_v1(q, q, q, q, 2, 1.0, torch.zeros(1, 1, dtype=torch.int32),
                               torch.zeros(1, dtype=torch.int32), 16, 1, None, "auto")
    with pytest.raises(RuntimeError, match="no CPU path"):
        cache_ops.reshape_and_cache(q, q, q, q, torch.zeros(1, dtype=torch.long), "auto")
    with pytest.raises(RuntimeError, match="Invalid device combination"):
        cache_ops.swap_blocks(q, q, {0: 0})


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "vllm_splitwise_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".inc", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", text, flags=re.M), f
                assert "oracle/" not in text and "liboracle" not in text, f
    code = ("import sys; sys.path.insert(0, %r); import vllm_splitwise_b200._C, vllm_splitwise_b200.paged_attn, "
            "vllm_splitwise_b200.cache_engine, vllm_splitwise_b200.comm_utils, vllm_splitwise_b200.sharding; "
            "assert not [m for m in sys.modules if m == 'oracle' or m.startswith('oracle.')]") % ROOT
    subprocess.run([sys.executable, "-c", code], check=True)


def test_missing_library_fails_loudly(tmp_path, monkeypatch):
    from vllm_splitwise_b200 import _lib
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", str(tmp_path / "libb200kv.so"))
    with pytest.raises(ImportError, match="no CPU fallback"):
        _lib.load()
