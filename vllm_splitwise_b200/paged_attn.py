"""Host-side op wrapper: mirror of the reference's PagedAttentionImpl
(vllm/model_executor/layers/attention/ops/paged_attn.py:15-111) over the B200 ops.

Same names, arguments and v1/v2 selection rule, so the attention backends' call sites
(backends/flash_attn.py:77-79,115-125, backends/xformers.py:178-188) work unchanged.
`input_metadata` is duck-typed: it needs slot_mapping, kv_cache_dtype, block_tables,
context_lens and max_context_len (vllm/model_executor/input_metadata.py).
"""
from typing import List, Optional

import torch

from ._C import cache_ops, ops

# Must equal the partition length the v2 kernels use (reference paged_attn.py:11-12).
_PARTITION_SIZE = 512


class PagedAttentionImpl:

    @staticmethod
    def get_supported_head_sizes() -> List[int]:
        return [64, 80, 96, 112, 128, 256]

    @staticmethod
    def reshape_and_cache(key: torch.Tensor, value: torch.Tensor, key_cache: torch.Tensor,
                          value_cache: torch.Tensor, input_metadata) -> None:
        cache_ops.reshape_and_cache(key, value, key_cache, value_cache,
                                    input_metadata.slot_mapping.flatten(),
                                    input_metadata.kv_cache_dtype)

    @staticmethod
    def use_v1(max_context_len: int, num_seqs: int, num_heads: int) -> bool:
        """The reference's selection rule (paged_attn.py:55-63), kept verbatim in meaning: v1 unless
        the context needs more than one 512-token partition AND the (seq, head) grid is small, or the
        context exceeds 8192."""
        max_num_partitions = (max_context_len + _PARTITION_SIZE - 1) // _PARTITION_SIZE
        return max_context_len <= 8192 and (max_num_partitions == 1 or num_seqs * num_heads > 512)

    @staticmethod
    def forward_decode(query: torch.Tensor, key_cache: torch.Tensor, value_cache: torch.Tensor,
                       input_metadata, num_kv_heads: int, scale: float,
                       alibi_slopes: Optional[torch.Tensor]) -> torch.Tensor:
        output = torch.empty_like(query)
        block_size = value_cache.shape[3]
        num_seqs, num_heads, head_size = query.shape
        max_context_len = input_metadata.max_context_len
        if PagedAttentionImpl.use_v1(max_context_len, num_seqs, num_heads):
            ops.paged_attention_v1(output, query, key_cache, value_cache, num_kv_heads, scale,
                                   input_metadata.block_tables, input_metadata.context_lens,
                                   block_size, max_context_len, alibi_slopes,
                                   input_metadata.kv_cache_dtype)
        else:
            assert _PARTITION_SIZE % block_size == 0
            max_num_partitions = (max_context_len + _PARTITION_SIZE - 1) // _PARTITION_SIZE
            tmp_output = torch.empty(size=(num_seqs, num_heads, max_num_partitions, head_size),
                                     dtype=output.dtype, device=output.device)
            exp_sums = torch.empty(size=(num_seqs, num_heads, max_num_partitions),
                                   dtype=torch.float32, device=output.device)
            max_logits = torch.empty_like(exp_sums)
            ops.paged_attention_v2(output, exp_sums, max_logits, tmp_output, query, key_cache,
                                   value_cache, num_kv_heads, scale, input_metadata.block_tables,
                                   input_metadata.context_lens, block_size, max_context_len,
                                   alibi_slopes, input_metadata.kv_cache_dtype)
        return output

    @staticmethod
    def forward_prefix(query: torch.Tensor, key: torch.Tensor, value: torch.Tensor, key_cache: torch.Tensor,
                       value_cache: torch.Tensor, input_metadata,
                       alibi_slopes: Optional[torch.Tensor]) -> torch.Tensor:
        """reference paged_attn.py:113-138; `input_metadata` needs block_tables, start_loc, prompt_lens,
        context_lens and max_seq_len."""
        from .prefix_prefill import context_attention_fwd
        output = torch.empty_like(query)
        context_attention_fwd(query, key, value, output, key_cache, value_cache,
                              input_metadata.block_tables,  # [BS, max_block_per_request]
                              input_metadata.start_loc, input_metadata.prompt_lens,
                              input_metadata.context_lens, input_metadata.max_seq_len, alibi_slopes)
        return output
