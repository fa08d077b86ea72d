"""Host-side helpers on the path (mirrors of the reference's vllm/utils.py entries it needs)."""
from typing import Dict, List, Optional, Tuple, Union

import torch

MAX_SLOT_IDS = 256  # semaphores / in-flight sequences for the KV transfer (reference vllm/utils.py:130)

STR_DTYPE_TO_TORCH_DTYPE = {
    "half": torch.half,
    "bfloat16": torch.bfloat16,
    "float": torch.float,
    "fp8_e5m2": torch.uint8,
}


def get_max_shared_memory_bytes(gpu: int = 0) -> int:
    """Max opt-in dynamic shared memory per block (reference vllm/utils.py get_max_shared_memory_bytes)."""
    from ._C import cuda_utils
    return int(cuda_utils.get_max_shared_memory_per_block_device_attribute(gpu))


def coalesce_blocks(block_list: List[int]) -> List[Tuple[int, int]]:
    """Merge block ids into (start, length) runs of consecutive ids -- the message list of the
    KV transfer (reference vllm/utils.py:403-420)."""
    runs: List[Tuple[int, int]] = []
    start = length = None
    for b in sorted(block_list):
        if start is not None and b == start + length:
            length += 1
            continue
        if start is not None:
            runs.append((start, length))
        start, length = b, 1
    if start is not None:
        runs.append((start, length))
    return runs


def coalesce_blocks_by_id(blocks_to_nw: Dict[int, List[int]]) -> Dict[int, List[Tuple[int, int]]]:
    """Per-semaphore version (reference vllm/utils.py:423-426)."""
    for key in blocks_to_nw:
        blocks_to_nw[key] = coalesce_blocks(blocks_to_nw[key])
    return blocks_to_nw


def kv_cache_shapes(num_blocks: int, block_size: int, num_kv_heads: int, head_size: int,
                    dtype: torch.dtype) -> Tuple[Tuple[int, ...], Tuple[int, ...]]:
    """Paged layouts (reference vllm/worker/cache_engine.py:60-75): key [NB,H,D/x,bs,x], value [NB,H,D,bs]."""
    x = 16 // torch.tensor([], dtype=dtype).element_size()
    return ((num_blocks, num_kv_heads, head_size // x, block_size, x),
            (num_blocks, num_kv_heads, head_size, block_size))


def _fill_random_fp8_e5m2(tensor: torch.Tensor, low: float, high: float) -> None:
    # uniform fp16 values pushed through the cache's own converter: never produces Inf/NaN bytes
    # (reference vllm/utils.py:311-329)
    from ._C import cache_ops
    tmp = torch.empty_like(tensor, dtype=torch.float16)
    tmp.uniform_(low, high)
    cache_ops.convert_fp8_e5m2(tmp, tensor)


def create_kv_caches_with_random(
    num_blocks: int,
    block_size: int,
    num_layers: int,
    num_heads: int,
    head_size: int,
    cache_dtype: Optional[Union[str, torch.dtype]],
    model_dtype: Optional[Union[str, torch.dtype]] = None,
    seed: Optional[int] = 0,
    device: Optional[str] = "cuda",
) -> Tuple[List[torch.Tensor], List[torch.Tensor]]:
    """Test/benchmark fixture with the reference's recipe (vllm/utils.py:332-397):
    uniform(-D^-0.5, D^-0.5) values in the paged layouts."""
    torch.random.manual_seed(seed)
    if torch.cuda.is_available():
        torch.cuda.manual_seed(seed)
    if isinstance(cache_dtype, str):
        if cache_dtype == "auto":
            if isinstance(model_dtype, str):
                torch_dtype = STR_DTYPE_TO_TORCH_DTYPE[model_dtype]
            elif isinstance(model_dtype, torch.dtype):
                torch_dtype = model_dtype
            else:
                raise ValueError(f"Invalid model dtype: {model_dtype}")
        elif cache_dtype in ("half", "bfloat16", "float", "fp8_e5m2"):
            torch_dtype = STR_DTYPE_TO_TORCH_DTYPE[cache_dtype]
        else:
            raise ValueError(f"Invalid kv cache dtype: {cache_dtype}")
    elif isinstance(cache_dtype, torch.dtype):
        torch_dtype = cache_dtype
    else:
        raise ValueError(f"Invalid kv cache dtype: {cache_dtype}")

    scale = head_size ** -0.5
    kshape, vshape = kv_cache_shapes(num_blocks, block_size, num_heads, head_size, torch_dtype)
    key_caches, value_caches = [], []
    for shape, bucket in ((kshape, key_caches), (vshape, value_caches)):
        for _ in range(num_layers):
            t = torch.empty(size=shape, dtype=torch_dtype, device=device)
            if cache_dtype == "fp8_e5m2":
                _fill_random_fp8_e5m2(t, -scale, scale)
            elif torch_dtype in (torch.half, torch.bfloat16, torch.float):
                t.uniform_(-scale, scale)
            else:
                raise ValueError(f"Does not support cache of type {cache_dtype}")
            bucket.append(t)
    return key_caches, value_caches
