"""Host-side mirror of the reference's prefix-prefill op
(vllm/model_executor/layers/attention/ops/prefix_prefill.py:621-745): same name, argument order and
meaning as `context_attention_fwd`, implemented by the sm_100a kernel behind b200_context_attention_fwd
(include/b200_kv.h).  The reference's is a Triton kernel; this path has no Triton and no fallback.
"""
import ctypes
from typing import Optional

import torch

from . import _lib
from ._C._common import device_ctx, dtype_code, require_cuda, stream_ptr


def _i32(t: torch.Tensor) -> torch.Tensor:
    # the reference's metadata is int64 in its test (test_prefix_prefill.py:66-72) and mixed int32/int64 in the
    # model runner; the kernel reads int32.  Prefill is not CUDA-graph captured (model_runner.py:593-598).
    return t if t.dtype == torch.int32 else t.to(torch.int32)


def context_attention_fwd(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, o: torch.Tensor,
                          k_cache: torch.Tensor, v_cache: torch.Tensor, b_loc: torch.Tensor,
                          b_start_loc: torch.Tensor, b_seq_len: torch.Tensor, b_ctx_len: torch.Tensor,
                          max_input_len: int, alibi_slopes: Optional[torch.Tensor] = None) -> None:
    for name, t in (("q", q), ("k", k), ("v", v), ("o", o), ("k_cache", k_cache), ("v_cache", v_cache),
                    ("b_loc", b_loc), ("b_start_loc", b_start_loc), ("b_seq_len", b_seq_len),
                    ("b_ctx_len", b_ctx_len)):
        require_cuda(t, name)
    Lq, Lk, Lv = q.shape[-1], k.shape[-1], v.shape[-1]
    assert Lq == Lk and Lk == Lv                       # reference :637-639
    assert Lk in {16, 32, 64, 128}
    if q.dtype not in (torch.float16, torch.bfloat16):
        raise RuntimeError(f"Unsupported data type: {q.dtype} (context_attention_fwd needs float16 or bfloat16)")
    for name, t in (("k", k), ("v", v), ("o", o), ("k_cache", k_cache), ("v_cache", v_cache)):
        if t.dtype != q.dtype:
            raise RuntimeError(f"Unsupported data type: {name} is {t.dtype}, query is {q.dtype}")
    for name, t in (("q", q), ("k", k), ("v", v), ("o", o)):
        if t.dim() != 3 or t.stride(2) != 1:
            raise RuntimeError(f"{name} must be [num_tokens, heads, head_size] with a dense last dim")
    bs = v_cache.shape[3]
    if (k_cache.dim() != 5 or k_cache.shape[4] != 8 or k_cache.shape[3] != bs or k_cache.stride(4) != 1
            or k_cache.stride(3) != 8 or k_cache.stride(2) != 8 * bs or v_cache.stride(3) != 1
            or v_cache.stride(2) != bs):
        raise RuntimeError("k_cache must be [num_blocks, kv_heads, head_size/8, block_size, 8] and v_cache "
                           "[num_blocks, kv_heads, head_size, block_size], dense inside a (block, head) chunk")
    b_loc, b_start_loc = _i32(b_loc), _i32(b_start_loc)
    b_seq_len, b_ctx_len = _i32(b_seq_len), _i32(b_ctx_len)
    if b_loc.stride(1) != 1:
        b_loc = b_loc.contiguous()
    a = _lib.ContextAttnArgs()
    a.out, a.query, a.key, a.value = o.data_ptr(), q.data_ptr(), k.data_ptr(), v.data_ptr()
    a.key_cache, a.value_cache = k_cache.data_ptr(), v_cache.data_ptr()
    a.block_tables = b_loc.data_ptr()
    a.start_loc = b_start_loc.contiguous().data_ptr()
    a.seq_lens = b_seq_len.contiguous().data_ptr()
    a.ctx_lens = b_ctx_len.contiguous().data_ptr()
    if alibi_slopes is not None:
        if alibi_slopes.dtype != torch.float32:
            raise RuntimeError("alibi_slopes must be float32")
        a.alibi_slopes = alibi_slopes.contiguous().data_ptr()
    a.scale = 1.0 / (Lq ** 0.5)                        # reference :641
    a.batch, a.num_heads, a.num_kv_heads = b_seq_len.shape[0], q.shape[1], k.shape[1]
    a.head_size, a.block_size = Lk, bs
    a.max_input_len = int(max_input_len)
    a.block_table_stride = b_loc.stride(0)
    a.q_stride_t, a.q_stride_h = q.stride(0), q.stride(1)
    a.k_stride_t, a.k_stride_h = k.stride(0), k.stride(1)
    a.v_stride_t, a.v_stride_h = v.stride(0), v.stride(1)
    a.o_stride_t, a.o_stride_h = o.stride(0), o.stride(1)
    a.kc_block_stride, a.kc_head_stride = k_cache.stride(0), k_cache.stride(1)
    a.vc_block_stride, a.vc_head_stride = v_cache.stride(0), v_cache.stride(1)
    a.dtype = dtype_code(q, "q")
    a.device = q.device.index if q.device.index is not None else -1
    with device_ctx(q.device):
        st = _lib.load().b200_context_attention_fwd(ctypes.byref(a), stream_ptr(q.device))
    _lib.check(st, "context_attention_fwd", a.head_size if st == _lib.ERR_HEAD_SIZE else
               a.block_size if st == _lib.ERR_BLOCK_SIZE else "")
