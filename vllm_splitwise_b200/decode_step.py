"""The composed decode step of the hot path, captured as a CUDA graph per padded batch size.

What the reference does per decode step and layer, and where:
    ModelRunner._prepare_decode      lists -> positions / slot_mapping / context_lens / block_tables, uploaded
                                     every step                          vllm/worker/model_runner.py:274-399
    LlamaAttention.forward           rotary_emb(positions, q, k)         vllm/model_executor/models/llama.py:156-158
    PagedAttentionImpl               reshape_and_cache, forward_decode   .../attention/ops/paged_attn.py:21-111
    CUDAGraphRunner                  copies the step's inputs into static buffers, graph.replay()
                                                                         vllm/worker/model_runner.py:826-873
Here the metadata lives on the device (`DeviceDecodeState`), RoPE is fused into the cache write, and the three
kernels of a layer go out as one `ops.decode_step` call whose launches are programmatically chained (PDL).
`DecodeStepRunner.capture()` records, for one batch size:
    H2D  one copy of the step's packed qkv rows from a pinned staging buffer      (inputs of the step)
    L x  [metadata (first layer only)] -> RoPE + cache write -> paged attention   (one ops.decode_step per layer)
    D2H  one copy of the attention output into a pinned buffer                    (result of the step)
so a step costs the host one graph launch and one stream synchronise.
"""
from typing import List, Optional, Sequence, Tuple

import torch

from ._C import ops
from .decode_metadata import DeviceDecodeState
from .paged_attn import _PARTITION_SIZE, PagedAttentionImpl


class CapturedDecodeStep:
    """One captured graph: static buffers + replay.  qkv_host is [layers, batch, (Hq + 2 Hkv) * D] pinned; fill it
    (or leave the previous contents) and call `replay()`; `out_host` [layers, batch, Hq, D] holds the result after
    `synchronize()`."""

    def __init__(self, graph, stream, qkv_host, qkv, out, out_host, batch) -> None:
        self.graph, self.stream = graph, stream
        self.qkv_host, self.qkv, self.out, self.out_host, self.batch = qkv_host, qkv, out, out_host, batch

    def replay(self) -> None:
        """Launch the step on torch's current stream."""
        self.graph.replay()

    def synchronize(self) -> None:
        torch.cuda.current_stream(self.out.device).synchronize()


class DecodeStepRunner:

    def __init__(self, state: DeviceDecodeState, num_heads: int, num_kv_heads: int, head_size: int, scale: float,
                 dtype: torch.dtype, kv_cache_dtype: str = "auto", cos_sin_cache: Optional[torch.Tensor] = None,
                 is_neox: bool = True, alibi_slopes: Optional[torch.Tensor] = None) -> None:
        self.state = state
        self.device = state.device
        self.num_heads, self.num_kv_heads, self.head_size = num_heads, num_kv_heads, head_size
        self.scale, self.dtype, self.kv_cache_dtype = float(scale), dtype, kv_cache_dtype
        self.cos_sin_cache, self.is_neox, self.alibi_slopes = cos_sin_cache, is_neox, alibi_slopes
        self.q_size = num_heads * head_size
        self.kv_size = num_kv_heads * head_size

    # ---- one layer, eager (also what capture() records) -------------------------------------------------
    def layer_step(self, qkv: torch.Tensor, out: torch.Tensor, key_cache: torch.Tensor, value_cache: torch.Tensor,
                   batch: int, max_context_len: int, with_metadata: bool, out_blocks: Optional[int] = None,
                   v2_scratch: Optional[Tuple[torch.Tensor, torch.Tensor, torch.Tensor]] = None) -> None:
        """qkv [batch, (Hq + 2 Hkv) * D] as qkv_proj leaves it; q, k, v are its split views (llama.py:155)."""
        st = self.state
        q, k, v = qkv.split([self.q_size, self.kv_size, self.kv_size], dim=-1)
        q = q.view(batch, self.num_heads, self.head_size)
        k = k.view(batch, self.num_kv_heads, self.head_size)
        v = v.view(batch, self.num_kv_heads, self.head_size)
        ob = out_blocks or st.max_blocks
        out_bt = st.out_block_tables.view(-1)[:batch * ob].view(batch, ob)
        es, ml, tmp = v2_scratch if v2_scratch is not None else (None, None, None)
        ops.decode_step(out, q, k, v, key_cache, value_cache, self.num_kv_heads, self.scale, out_bt,
                        st.context_lens[:batch], value_cache.shape[3], max_context_len, self.alibi_slopes,
                        self.kv_cache_dtype, st.positions.view(-1)[:batch], st.slot_mapping.view(-1)[:batch],
                        self.cos_sin_cache, self.is_neox,
                        st.block_tables if with_metadata else None, st.seq_lens if with_metadata else None,
                        st.batch_rows if with_metadata else None, st.sliding_window, es, ml, tmp)

    def v2_scratch(self, batch: int, max_context_len: int):
        """The reference's v2 scratch (paged_attn.py:83-93) when its rule picks v2, else None."""
        if PagedAttentionImpl.use_v1(max_context_len, batch, self.num_heads):
            return None
        parts = (max_context_len + _PARTITION_SIZE - 1) // _PARTITION_SIZE
        tmp = torch.empty(batch, self.num_heads, parts, self.head_size, dtype=self.dtype, device=self.device)
        es = torch.empty(batch, self.num_heads, parts, dtype=torch.float32, device=self.device)
        return es, torch.empty_like(es), tmp

    # ---- capture ------------------------------------------------------------------------------------
    def capture(self, batch: int, kv_caches: Sequence[Tuple[torch.Tensor, torch.Tensor]], max_context_len: int,
                out_blocks: Optional[int] = None, warmup: int = 2) -> CapturedDecodeStep:
        """Record one decode step over `kv_caches` (one (key_cache, value_cache) per layer) for `batch` rows of
        the state's current batch (`state.set_batch`).  max_context_len is baked in, as the reference bakes
        max_context_len_to_capture (model_runner.py:751); the kernels bound their work by context_lens."""
        L = len(kv_caches)
        width = self.q_size + 2 * self.kv_size
        qkv_host = torch.zeros(L, batch, width, dtype=self.dtype).pin_memory()
        out_host = torch.zeros(L, batch, self.num_heads, self.head_size, dtype=self.dtype).pin_memory()
        qkv = torch.zeros(L, batch, width, dtype=self.dtype, device=self.device)
        out = torch.zeros(L, batch, self.num_heads, self.head_size, dtype=self.dtype, device=self.device)
        scratch = self.v2_scratch(batch, max_context_len)
        stream = torch.cuda.Stream(self.device)

        def body():
            qkv.copy_(qkv_host, non_blocking=True)
            for layer, (kc, vc) in enumerate(kv_caches):
                self.layer_step(qkv[layer], out[layer], kc, vc, batch, max_context_len, with_metadata=layer == 0,
                                out_blocks=out_blocks, v2_scratch=scratch)
            out_host.copy_(out, non_blocking=True)

        # Warm-up runs outside the capture (function attributes, workspace allocation) must not touch the
        # sequences: they run with every batch entry marked as padding (row -1: nothing advances, no cache write).
        rows = self.state.batch_rows[:batch]
        stream.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(stream):
            saved = rows.clone()
            rows.fill_(-1)
            for _ in range(warmup):
                body()
            rows.copy_(saved)
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=stream):
                body()
        stream.synchronize()
        torch.cuda.current_stream(self.device).wait_stream(stream)
        return CapturedDecodeStep(graph, stream, qkv_host, qkv, out, out_host, batch)
