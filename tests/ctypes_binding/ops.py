"""TEST HELPER -- a second, independent binding of the C ABI (ctypes instead of pybind11), used by the tests to
prove that include/b200_kv.h is bindable without torch types and that both bindings give identical bits.
The product binding is vllm_splitwise_b200/csrc/torch_binding.cpp.

vllm._C.ops on the hot path: paged_attention_v1 / paged_attention_v2.

Positional signatures are the reference's (csrc/ops.h:5-34); all tensors are caller-allocated,
ops write in place and return None, launches go to torch's current stream.
"""
import ctypes
from typing import Optional

import torch

from vllm_splitwise_b200 import _lib
from vllm_splitwise_b200._C._common import device_ctx, dtype_code, require_cuda, stream_ptr



def _cache_dtype_code(kv_cache_dtype: str, query: torch.Tensor, cache: torch.Tensor) -> int:
    if kv_cache_dtype == "auto":
        if cache.dtype != query.dtype:
            raise RuntimeError(f"Unsupported data type: cache {cache.dtype} with query {query.dtype}")
        return dtype_code(query, "query")
    if kv_cache_dtype == "fp8_e5m2":
        if cache.dtype != torch.uint8:
            raise RuntimeError(f"Unsupported data type: fp8_e5m2 cache must be uint8, got {cache.dtype}")
        return _lib.U8_E5M2
    raise RuntimeError(f"Unsupported data type of kv cache: {kv_cache_dtype}")


def _fill_args(out, query, key_cache, value_cache, num_kv_heads, scale, block_tables, context_lens,
               block_size, max_context_len, alibi_slopes, kv_cache_dtype) -> _lib.PagedAttnArgs:
    for name, t in (("out", out), ("query", query), ("key_cache", key_cache),
                    ("value_cache", value_cache), ("block_tables", block_tables),
                    ("context_lens", context_lens)):
        require_cuda(t, name)
    if block_tables.dtype != torch.int32 or context_lens.dtype != torch.int32:
        raise RuntimeError("block_tables and context_lens must be int32")
    if query.stride(-1) != 1 or query.stride(1) != query.size(2) or not out.is_contiguous():
        raise RuntimeError("query must be dense in its last two dims and out contiguous")
    a = _lib.PagedAttnArgs()
    a.out = out.data_ptr()
    a.query = query.data_ptr()
    a.key_cache = key_cache.data_ptr()
    a.value_cache = value_cache.data_ptr()
    a.block_tables = block_tables.data_ptr()
    a.context_lens = context_lens.data_ptr()
    if alibi_slopes is not None:
        if alibi_slopes.dtype != torch.float32:
            raise RuntimeError("alibi_slopes must be float32")
        a.alibi_slopes = alibi_slopes.data_ptr()
    a.scale = float(scale)
    a.num_seqs, a.num_heads, a.head_size = query.size(0), query.size(1), query.size(2)
    a.num_kv_heads = int(num_kv_heads)
    a.block_size = int(block_size)
    a.max_context_len = int(max_context_len)
    a.max_num_blocks_per_seq = block_tables.size(1)
    a.q_stride = query.stride(0)
    a.kv_block_stride = key_cache.stride(0)
    a.kv_head_stride = key_cache.stride(1)
    a.query_dtype = dtype_code(query, "query")
    a.cache_dtype = _cache_dtype_code(kv_cache_dtype, query, key_cache)
    a.device = query.device.index if query.device.index is not None else -1
    return a


def _detail(a: _lib.PagedAttnArgs, status: int):
    if status == _lib.ERR_HEAD_SIZE:
        return a.head_size
    if status == _lib.ERR_BLOCK_SIZE:
        return a.block_size
    return ""


def _workspace_for(a: _lib.PagedAttnArgs, device: torch.device) -> Optional[torch.Tensor]:
    """Test helper: a fresh zeroed buffer of exactly the size this call needs (never shared between calls;
    the product binding, csrc/torch_binding.cpp, keeps one per stream and one per captured graph)."""
    need = _lib.load().b200_paged_attention_workspace_bytes(ctypes.byref(a))
    if need <= 0:
        return None
    return torch.zeros(need, dtype=torch.uint8, device=device)


def paged_attention_v1(out: torch.Tensor, query: torch.Tensor, key_cache: torch.Tensor,
                       value_cache: torch.Tensor, num_kv_heads: int, scale: float,
                       block_tables: torch.Tensor, context_lens: torch.Tensor, block_size: int,
                       max_context_len: int, alibi_slopes: Optional[torch.Tensor],
                       kv_cache_dtype: str) -> None:
    a = _fill_args(out, query, key_cache, value_cache, num_kv_heads, scale, block_tables,
                   context_lens, block_size, max_context_len, alibi_slopes, kv_cache_dtype)
    # Scratch that lets the library split long contexts into partitions so that the persistent grid's last
    # round is full whatever the batch / head count is (partials are combined inside the same kernel).
    if a.max_context_len >= 256:
        ws = _workspace_for(a, query.device)
        if ws is not None:
            a.workspace = ws.data_ptr()
            a.workspace_bytes = ws.numel()
    st = _lib.load().b200_paged_attention_v1(ctypes.byref(a), stream_ptr(query.device))
    _lib.check(st, "paged_attention_v1", _detail(a, st))


def paged_attention_v2(out: torch.Tensor, exp_sums: torch.Tensor, max_logits: torch.Tensor,
                       tmp_out: torch.Tensor, query: torch.Tensor, key_cache: torch.Tensor,
                       value_cache: torch.Tensor, num_kv_heads: int, scale: float,
                       block_tables: torch.Tensor, context_lens: torch.Tensor, block_size: int,
                       max_context_len: int, alibi_slopes: Optional[torch.Tensor],
                       kv_cache_dtype: str) -> None:
    a = _fill_args(out, query, key_cache, value_cache, num_kv_heads, scale, block_tables,
                   context_lens, block_size, max_context_len, alibi_slopes, kv_cache_dtype)
    if exp_sums.dtype != torch.float32 or max_logits.dtype != torch.float32:
        raise RuntimeError("exp_sums and max_logits must be float32")
    if tmp_out.dtype != query.dtype:
        raise RuntimeError("tmp_out must have the query dtype")
    if not (exp_sums.is_contiguous() and max_logits.is_contiguous() and tmp_out.is_contiguous()):
        raise RuntimeError("v2 scratch tensors must be contiguous")
    a.exp_sums = exp_sums.data_ptr()
    a.max_logits = max_logits.data_ptr()
    a.tmp_out = tmp_out.data_ptr()
    a.max_num_partitions = exp_sums.size(2)
    st = _lib.load().b200_paged_attention_v2(ctypes.byref(a), stream_ptr(query.device))
    _lib.check(st, "paged_attention_v2", _detail(a, st))


def _rotary_shape(positions, query, key, head_size, cos_sin_cache):
    for name, t in (("positions", positions), ("query", query), ("key", key), ("cos_sin_cache", cos_sin_cache)):
        require_cuda(t, name)
    if positions.dtype != torch.int64 or not positions.is_contiguous():
        raise RuntimeError("positions must be a contiguous int64 tensor")
    if (query.dim() < 2 or key.dim() < 2 or head_size <= 0 or query.size(-1) % head_size or key.size(-1) % head_size
            or query.stride(-1) != 1 or key.stride(-1) != 1):
        raise RuntimeError("query/key must be [..., num_heads * head_size] with a dense last dim")
    if key.dtype != query.dtype or cos_sin_cache.dtype != query.dtype:
        raise RuntimeError("query, key and cos_sin_cache must have the same dtype")
    if cos_sin_cache.dim() != 2 or not cos_sin_cache.is_contiguous():
        raise RuntimeError("cos_sin_cache must be [max_position, rot_dim]")
    num_tokens = query.numel() // query.size(-1)
    if positions.numel() != num_tokens or key.numel() // key.size(-1) != num_tokens:
        raise RuntimeError("positions, query and key must describe the same number of tokens")
    return (num_tokens, query.size(-1) // head_size, key.size(-1) // head_size, cos_sin_cache.size(1),
            query.stride(-2), key.stride(-2))


def rotary_embedding(positions: torch.Tensor, query: torch.Tensor, key: torch.Tensor, head_size: int,
                     cos_sin_cache: torch.Tensor, is_neox: bool) -> None:
    """reference csrc/ops.h:48-54"""
    n, hq, hkv, rot, qs, ks = _rotary_shape(positions, query, key, head_size, cos_sin_cache)
    with device_ctx(query.device):
        st = _lib.load().b200_rotary_embedding(positions.data_ptr(), query.data_ptr(), key.data_ptr(),
                                               cos_sin_cache.data_ptr(), None, n, hq, hkv, head_size, rot, qs, ks,
                                               int(bool(is_neox)), dtype_code(query, "query"), stream_ptr(query.device))
    _lib.check(st, "rotary_embedding")


def batched_rotary_embedding(positions: torch.Tensor, query: torch.Tensor, key: torch.Tensor, head_size: int,
                             cos_sin_cache: torch.Tensor, is_neox: bool, rot_dim: int,
                             cos_sin_cache_offsets: torch.Tensor) -> None:
    """reference csrc/ops.h:56-63"""
    n, hq, hkv, rot, qs, ks = _rotary_shape(positions, query, key, head_size, cos_sin_cache)
    require_cuda(cos_sin_cache_offsets, "cos_sin_cache_offsets")
    if (cos_sin_cache_offsets.dtype != torch.int64 or not cos_sin_cache_offsets.is_contiguous()
            or cos_sin_cache_offsets.numel() != n):
        raise RuntimeError("cos_sin_cache_offsets must be a contiguous int64 tensor with one entry per token")
    if rot_dim != rot:
        raise RuntimeError("rot_dim must equal cos_sin_cache.size(1)")
    with device_ctx(query.device):
        st = _lib.load().b200_rotary_embedding(positions.data_ptr(), query.data_ptr(), key.data_ptr(),
                                               cos_sin_cache.data_ptr(), cos_sin_cache_offsets.data_ptr(), n, hq, hkv,
                                               head_size, rot, qs, ks, int(bool(is_neox)), dtype_code(query, "query"),
                                               stream_ptr(query.device))
    _lib.check(st, "batched_rotary_embedding")
