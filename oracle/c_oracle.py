"""TEST INFRASTRUCTURE ONLY -- ctypes front of oracle/_build/liboracle_c.so (oracle/oracle_c.c)."""
import ctypes
import os
from ctypes import POINTER, c_float, c_int, c_int32, c_int64, c_uint8, c_void_p

import numpy as np
import torch

_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_build", "liboracle_c.so")
_lib = None


def load():
    global _lib
    if _lib is None:
        if not os.path.exists(_PATH):
            raise ImportError(f"{_PATH} missing: run oracle/build_oracle.sh (or __graft_entry__.build())")
        _lib = ctypes.CDLL(_PATH)
        _lib.oracle_e5m2_encode.argtypes = [c_void_p, c_void_p, c_int64]
        _lib.oracle_e5m2_decode.argtypes = [c_void_p, c_void_p, c_int64]
        _lib.oracle_reshape_and_cache.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int,
                                                  c_int, c_int, c_int, c_int64, c_int64, c_int]
        _lib.oracle_copy_blocks.argtypes = [c_void_p, c_void_p, c_int, c_int64]
        _lib.oracle_paged_attention_f32.argtypes = [c_void_p] * 7 + [c_float] + [c_int] * 7
    return _lib


def e5m2_encode(x: np.ndarray) -> np.ndarray:
    x = np.ascontiguousarray(x, dtype=np.float32)
    out = np.empty(x.shape, dtype=np.uint8)
    load().oracle_e5m2_encode(x.ctypes.data, out.ctypes.data, x.size)
    return out


def e5m2_decode(b: np.ndarray) -> np.ndarray:
    b = np.ascontiguousarray(b, dtype=np.uint8)
    out = np.empty(b.shape, dtype=np.float32)
    load().oracle_e5m2_decode(b.ctypes.data, out.ctypes.data, b.size)
    return out


def reshape_and_cache(key, value, key_cache, value_cache, slot_mapping) -> None:
    """Byte-exact scatter on CPU torch tensors ("auto" cache dtype); key/value may be strided views."""
    n, h, d = key.shape
    bs, x = key_cache.shape[3], key_cache.shape[4]
    assert key.stride(2) == 1 and key.stride(1) == d and value.stride(2) == 1 and value.stride(1) == d
    sm = slot_mapping.contiguous()
    load().oracle_reshape_and_cache(key.data_ptr(), value.data_ptr(), key_cache.data_ptr(), value_cache.data_ptr(),
                                    sm.data_ptr(), n, h, d, bs, x, key.stride(0), value.stride(0),
                                    key.element_size())


def copy_blocks(cache: torch.Tensor, pairs) -> None:
    arr = np.asarray(pairs, dtype=np.int64).reshape(-1)
    load().oracle_copy_blocks(cache.data_ptr(), arr.ctypes.data, arr.size // 2,
                              cache[0].numel() * cache.element_size())


def paged_attention(query, key_cache, value_cache, block_tables, context_lens, scale, alibi_slopes=None):
    """fp32 evaluation (inputs widened); returns float32 [S,H,D]."""
    q = query.float().contiguous()
    kc = key_cache.float().contiguous()
    vc = value_cache.float().contiguous()
    bt = block_tables.to(torch.int32).contiguous()
    cl = context_lens.to(torch.int32).contiguous()
    out = torch.empty_like(q)
    al = alibi_slopes.float().contiguous() if alibi_slopes is not None else None
    s, h, d = q.shape
    load().oracle_paged_attention_f32(out.data_ptr(), q.data_ptr(), kc.data_ptr(), vc.data_ptr(), bt.data_ptr(),
                                      cl.data_ptr(), al.data_ptr() if al is not None else None, float(scale), s, h,
                                      kc.shape[1], d, kc.shape[3], kc.shape[4], bt.shape[1])
    return out
