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


class PipelinedDecodeStep:
    """The same composed step with the copy engines overlapped: step i's packed qkv rows travel host -> device on
    one copy stream and step i's output device -> host on another while the kernels of neighbouring steps run,
    `depth` steps in flight (double buffering by default).  Per step the host issues two async copies and one
    kernel-only graph launch and waits for the result of an EARLIER step, never for the one it just submitted.

        p = PipelinedDecodeStep(runner, batch, kv_caches, max_context_len)
        p.slot(n).qkv_host[...] = step n's qkv rows;  p.submit(n, layer);  ...;  out = p.result(n - 1)

    Ordering per slot s = n % depth:  H2D(n) after the kernels of step n - depth (they read qkv[s]);
    kernels(n) after H2D(n) and after D2H(n - depth) (they overwrite out[s]);  D2H(n) after kernels(n).
    The kernels of consecutive steps stay in issue order on one compute stream: they share the decode state."""

    class _Slot:
        pass

    def __init__(self, runner: DecodeStepRunner, batch: int, kv_caches: Sequence[Tuple[torch.Tensor, torch.Tensor]],
                 max_context_len: int, depth: int = 2, out_blocks: Optional[int] = None) -> None:
        self.runner, self.batch, self.depth = runner, batch, depth
        dev = runner.device
        self.compute, self.h2d, self.d2h = torch.cuda.Stream(dev), torch.cuda.Stream(dev), torch.cuda.Stream(dev)
        width = runner.q_size + 2 * runner.kv_size
        scratch = runner.v2_scratch(batch, max_context_len)
        self.slots = []
        for _ in range(depth):
            s = self._Slot()
            s.qkv_host = torch.zeros(batch, width, dtype=runner.dtype).pin_memory()
            s.out_host = torch.zeros(batch, runner.num_heads, runner.head_size, dtype=runner.dtype).pin_memory()
            s.qkv = torch.zeros(batch, width, dtype=runner.dtype, device=dev)
            s.out = torch.zeros(batch, runner.num_heads, runner.head_size, dtype=runner.dtype, device=dev)
            s.ev_in, s.ev_done, s.ev_out = torch.cuda.Event(), torch.cuda.Event(), torch.cuda.Event()
            s.graphs = []
            self.slots.append(s)
        rows = runner.state.batch_rows[:batch]
        self.compute.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(self.compute):
            saved = rows.clone()
            rows.fill_(-1)  # warm-up as padding entries: nothing advances, no cache write (see capture())
            for s in self.slots:
                for kc, vc in kv_caches:
                    runner.layer_step(s.qkv, s.out, kc, vc, batch, max_context_len, True, out_blocks, scratch)
            rows.copy_(saved)
            for s in self.slots:
                for kc, vc in kv_caches:
                    g = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(g, stream=self.compute):
                        runner.layer_step(s.qkv, s.out, kc, vc, batch, max_context_len, True, out_blocks, scratch)
                    s.graphs.append(g)
        self.compute.synchronize()
        for s in self.slots:  # "already happened" for the first use of every slot
            s.ev_done.record(self.compute)
            s.ev_out.record(self.d2h)
        torch.cuda.current_stream(dev).wait_stream(self.compute)

    def slot(self, n: int):
        return self.slots[n % self.depth]

    def submit(self, n: int, layer: int = 0) -> None:
        s = self.slots[n % self.depth]
        with torch.cuda.stream(self.h2d):
            self.h2d.wait_event(s.ev_done)
            s.qkv.copy_(s.qkv_host, non_blocking=True)
            s.ev_in.record(self.h2d)
        with torch.cuda.stream(self.compute):
            self.compute.wait_event(s.ev_in)
            self.compute.wait_event(s.ev_out)
            s.graphs[layer].replay()
            s.ev_done.record(self.compute)
        with torch.cuda.stream(self.d2h):
            self.d2h.wait_event(s.ev_done)
            s.out_host.copy_(s.out, non_blocking=True)
            s.ev_out.record(self.d2h)

    def result(self, n: int) -> torch.Tensor:
        """Block until step n's output is in pinned host memory and return it (valid until step n + depth is
        submitted)."""
        s = self.slots[n % self.depth]
        s.ev_out.synchronize()
        return s.out_host
