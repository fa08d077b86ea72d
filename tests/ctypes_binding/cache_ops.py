"""vllm._C.cache_ops: reshape_and_cache, copy_blocks, swap_blocks, convert_fp8_e5m2
(reference csrc/cache.h:8-29, csrc/cache_kernels.cu)."""
import ctypes
from typing import Dict, List

import torch

from vllm_splitwise_b200 import _lib
from vllm_splitwise_b200._C._common import device_ctx, dtype_code, require_cuda, stream_ptr


def reshape_and_cache(key: torch.Tensor, value: torch.Tensor, key_cache: torch.Tensor,
                      value_cache: torch.Tensor, slot_mapping: torch.Tensor,
                      kv_cache_dtype: str) -> None:
    for name, t in (("key", key), ("value", value), ("key_cache", key_cache),
                    ("value_cache", value_cache), ("slot_mapping", slot_mapping)):
        require_cuda(t, name)
    if slot_mapping.dtype != torch.int64 or not slot_mapping.is_contiguous():
        raise RuntimeError("slot_mapping must be a contiguous int64 tensor")
    if kv_cache_dtype == "auto":
        cache_code = dtype_code(key, "key")
        if key_cache.dtype != key.dtype:
            raise RuntimeError("Unsupported data type: cache dtype must equal key dtype for 'auto'")
    elif kv_cache_dtype == "fp8_e5m2":
        cache_code = _lib.U8_E5M2
        if key_cache.dtype != torch.uint8:
            raise RuntimeError("Unsupported data type: fp8_e5m2 cache must be uint8")
    else:
        raise RuntimeError(f"Unsupported data type of kv cache: {kv_cache_dtype}")
    for t in (key, value):
        if t.stride(-1) != 1 or t.stride(1) != t.size(2):
            raise RuntimeError("key/value must be dense in their last two dims")
    if not (key_cache.is_contiguous() and value_cache.is_contiguous()):
        raise RuntimeError("key_cache/value_cache must be contiguous")
    num_tokens, num_heads, head_size = key.shape
    block_size, x = key_cache.size(3), key_cache.size(4)
    with device_ctx(key.device):
        st = _lib.load().b200_reshape_and_cache(
            key.data_ptr(), value.data_ptr(), key_cache.data_ptr(), value_cache.data_ptr(),
            slot_mapping.data_ptr(), num_tokens, num_heads, head_size, block_size, x,
            key.stride(0), value.stride(0), key_cache.size(0), dtype_code(key, "key"), cache_code,
            stream_ptr(key.device))
    _lib.check(st, "reshape_and_cache")


def copy_blocks(key_caches: List[torch.Tensor], value_caches: List[torch.Tensor],
                block_mapping: Dict[int, List[int]]) -> None:
    num_layers = len(key_caches)
    if num_layers != len(value_caches):
        raise RuntimeError("key_caches and value_caches must have the same length")
    if num_layers == 0:
        return
    device = key_caches[0].device
    require_cuda(key_caches[0], "key_caches[0]")
    pairs: List[int] = []
    for src in sorted(block_mapping):  # std::map iteration order (cache_kernels.cu:119-126)
        for dst in block_mapping[src]:
            pairs += (int(src), int(dst))
    num_pairs = len(pairs) // 2
    if num_pairs == 0:
        return
    kp = (ctypes.c_int64 * num_layers)(*[t.data_ptr() for t in key_caches])
    vp = (ctypes.c_int64 * num_layers)(*[t.data_ptr() for t in value_caches])
    pa = (ctypes.c_int64 * len(pairs))(*pairs)
    block_bytes = key_caches[0][0].numel() * key_caches[0].element_size()
    with device_ctx(device):
        st = _lib.load().b200_copy_blocks(kp, vp, num_layers, pa, num_pairs, block_bytes,
                                          stream_ptr(device))
    _lib.check(st, "copy_blocks")


def swap_blocks(src: torch.Tensor, dst: torch.Tensor, block_mapping: Dict[int, int]) -> None:
    if src.is_cuda and dst.is_cuda:
        if src.device.index != dst.device.index:
            raise RuntimeError("src and dst must be on the same GPU")
        kind, device = _lib.SWAP_D2D, src.device
    elif src.is_cuda and not dst.is_cuda:
        kind, device = _lib.SWAP_D2H, src.device
    elif dst.is_cuda and not src.is_cuda:
        kind, device = _lib.SWAP_H2D, dst.device
    else:
        raise RuntimeError("Invalid device combination")
    pairs: List[int] = []
    for s in sorted(block_mapping):
        pairs += (int(s), int(block_mapping[s]))
    num_pairs = len(pairs) // 2
    if num_pairs == 0:
        return
    pa = (ctypes.c_int64 * len(pairs))(*pairs)
    block_bytes = src.element_size() * src[0].numel()
    with device_ctx(device):
        st = _lib.load().b200_swap_blocks(
            src.data_ptr(), dst.data_ptr(), block_bytes, pa, num_pairs, kind,
            src.device.index if src.is_cuda else -1, dst.device.index if dst.is_cuda else -1,
            stream_ptr(device))
    _lib.check(st, "swap_blocks")


def convert_fp8_e5m2(src_cache: torch.Tensor, dst_cache: torch.Tensor) -> None:
    require_cuda(src_cache, "src_cache")
    require_cuda(dst_cache, "dst_cache")
    if src_cache.numel() != dst_cache.numel():
        raise RuntimeError("src_cache and dst_cache must have the same number of elements")
    if not (src_cache.is_contiguous() and dst_cache.is_contiguous()):
        raise RuntimeError("convert_fp8_e5m2 needs contiguous tensors")
    with device_ctx(src_cache.device):
        st = _lib.load().b200_convert_fp8_e5m2(
            src_cache.data_ptr(), dst_cache.data_ptr(), src_cache.numel(),
            dtype_code(src_cache, "src_cache"), dtype_code(dst_cache, "dst_cache"),
            stream_ptr(src_cache.device))
    _lib.check(st, "convert_fp8_e5m2")


def rotary_embedding_reshape_and_cache(positions: torch.Tensor, query: torch.Tensor, key: torch.Tensor,
                                       value: torch.Tensor, head_size: int, cos_sin_cache: torch.Tensor,
                                       is_neox: bool, key_cache: torch.Tensor, value_cache: torch.Tensor,
                                       slot_mapping: torch.Tensor, kv_cache_dtype: str) -> None:
    """ops.rotary_embedding(positions, query, key, ...) followed by reshape_and_cache(key, value, ...) in one
    call; query/key/value are [..., heads * head_size]."""
    from .ops import _rotary_shape
    n, hq, hkv, rot, qs, ks = _rotary_shape(positions, query, key, head_size, cos_sin_cache)
    for name, t in (("value", value), ("key_cache", key_cache), ("value_cache", value_cache),
                    ("slot_mapping", slot_mapping)):
        require_cuda(t, name)
    if (value.dim() != key.dim() or value.size(-1) != key.size(-1) or value.stride(-1) != 1
            or value.dtype != key.dtype or value.numel() != key.numel()):
        raise RuntimeError("value must have the shape and dtype of key")
    if slot_mapping.dtype != torch.int64 or not slot_mapping.is_contiguous() or slot_mapping.numel() != n:
        raise RuntimeError("slot_mapping must be a contiguous int64 tensor with one entry per token")
    if not (key_cache.is_contiguous() and value_cache.is_contiguous()):
        raise RuntimeError("key_cache/value_cache must be contiguous")
    if key_cache.dim() != 5 or key_cache.size(1) != hkv:
        raise RuntimeError("key_cache must be [blocks, kv_heads, D/x, bs, x]")
    if kv_cache_dtype == "auto":
        cache_code = dtype_code(key, "key")
        if key_cache.dtype != key.dtype:
            raise RuntimeError("Unsupported data type: cache dtype must equal key dtype for 'auto'")
    elif kv_cache_dtype == "fp8_e5m2":
        cache_code = _lib.U8_E5M2
        if key_cache.dtype != torch.uint8:
            raise RuntimeError("Unsupported data type: fp8_e5m2 cache must be uint8")
    else:
        raise RuntimeError(f"Unsupported data type of kv cache: {kv_cache_dtype}")
    with device_ctx(query.device):
        st = _lib.load().b200_rotary_embedding_reshape_and_cache(
            positions.data_ptr(), query.data_ptr(), key.data_ptr(), value.data_ptr(), cos_sin_cache.data_ptr(), n, hq,
            hkv, head_size, rot, qs, ks, value.stride(-2), int(bool(is_neox)), key_cache.data_ptr(),
            value_cache.data_ptr(), slot_mapping.data_ptr(), key_cache.size(3), key_cache.size(4), key_cache.size(0),
            dtype_code(key, "key"), cache_code, stream_ptr(query.device))
    _lib.check(st, "rotary_embedding_reshape_and_cache")
