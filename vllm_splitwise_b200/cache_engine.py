"""KV-cache owner: mirror of the reference's CacheEngine (vllm/worker/cache_engine.py:15-174),
reduced to what the hot path needs -- the layouts, the per-layer (K, V) allocation, and the
swap_in / swap_out / copy entry points over the B200 cache ops.  It takes plain numbers instead of
the reference's config objects (those belong to the out-of-scope control plane).
"""
from typing import Dict, List, Tuple

import torch

from .utils import STR_DTYPE_TO_TORCH_DTYPE, kv_cache_shapes

KVCache = Tuple[torch.Tensor, torch.Tensor]


class CacheEngine:

    def __init__(self, *, num_layers: int, num_kv_heads: int, head_size: int, block_size: int,
                 num_gpu_blocks: int, num_cpu_blocks: int, model_dtype: torch.dtype,
                 cache_dtype: str = "auto", device: str = "cuda", pin_memory: bool = True) -> None:
        self.num_layers = num_layers
        self.num_heads = num_kv_heads
        self.head_size = head_size
        self.block_size = block_size
        self.num_gpu_blocks = num_gpu_blocks
        self.num_cpu_blocks = num_cpu_blocks
        self.device = device
        self.pin_memory = pin_memory
        self.cache_dtype = cache_dtype
        self.dtype = model_dtype if cache_dtype == "auto" else STR_DTYPE_TO_TORCH_DTYPE[cache_dtype]
        self.gpu_cache = self._allocate(num_gpu_blocks, device, False)
        self.cpu_cache = self._allocate(num_cpu_blocks, "cpu", pin_memory)
        # swaps run on their own stream, one event per layer (reference cache_engine.py:51-58,129-139)
        self.cache_stream = torch.cuda.Stream(device=device)
        self.events = [torch.cuda.Event() for _ in range(num_layers)]

    def get_key_block_shape(self) -> Tuple[int, int, int, int]:
        return kv_cache_shapes(1, self.block_size, self.num_heads, self.head_size, self.dtype)[0][1:]

    def get_value_block_shape(self) -> Tuple[int, int, int]:
        return kv_cache_shapes(1, self.block_size, self.num_heads, self.head_size, self.dtype)[1][1:]

    def _allocate(self, num_blocks: int, device: str, pin: bool) -> List[KVCache]:
        kshape, vshape = kv_cache_shapes(num_blocks, self.block_size, self.num_heads, self.head_size, self.dtype)
        kw = dict(dtype=self.dtype, device=device)
        if device == "cpu":
            kw["pin_memory"] = pin
        return [(torch.empty(kshape, **kw), torch.empty(vshape, **kw)) for _ in range(self.num_layers)]

    def _swap(self, src: List[KVCache], dst: List[KVCache], src_to_dst: Dict[int, int]) -> None:
        from ._C import cache_ops
        with torch.cuda.stream(self.cache_stream):
            for i in range(self.num_layers):
                cache_ops.swap_blocks(src[i][0], dst[i][0], src_to_dst)
                cache_ops.swap_blocks(src[i][1], dst[i][1], src_to_dst)
                self.events[i].record(stream=self.cache_stream)

    def swap_in(self, src_to_dst: Dict[int, int]) -> None:
        self._swap(self.cpu_cache, self.gpu_cache, src_to_dst)

    def swap_out(self, src_to_dst: Dict[int, int]) -> None:
        self._swap(self.gpu_cache, self.cpu_cache, src_to_dst)

    def copy(self, src_to_dsts: Dict[int, List[int]]) -> None:
        from ._C import cache_ops
        # one launch per 64 (layer, K|V) buffers, tables in kernel parameters: no host sync
        cache_ops.copy_blocks([k for k, _ in self.gpu_cache], [v for _, v in self.gpu_cache], src_to_dsts)

    @staticmethod
    def get_cache_block_size(block_size: int, cache_dtype: str, model_dtype: torch.dtype, head_size: int,
                             num_kv_heads: int, num_layers: int) -> int:
        """Bytes of one block across all layers, K+V (reference cache_engine.py:155-174)."""
        dtype = model_dtype if cache_dtype == "auto" else STR_DTYPE_TO_TORCH_DTYPE[cache_dtype]
        elt = torch.tensor([], dtype=dtype).element_size()
        return elt * num_layers * 2 * block_size * num_kv_heads * head_size
