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


def resolve_cache_dtype(cache_dtype, model_dtype=None) -> torch.dtype:
    """Storage dtype of a KV cache: "auto" follows the model dtype, "fp8_e5m2" is stored as uint8
    (reference vllm/utils.py:30-35); a torch.dtype passes through."""
    if isinstance(cache_dtype, torch.dtype):
        return cache_dtype
    if cache_dtype == "auto":
        cache_dtype = model_dtype
        if isinstance(cache_dtype, torch.dtype):
            return cache_dtype
        if cache_dtype not in STR_DTYPE_TO_TORCH_DTYPE:
            raise ValueError(f"Invalid model dtype: {model_dtype}")
    if not isinstance(cache_dtype, str) or cache_dtype not in STR_DTYPE_TO_TORCH_DTYPE:
        raise ValueError(f"Invalid kv cache dtype: {cache_dtype}")
    return STR_DTYPE_TO_TORCH_DTYPE[cache_dtype]


def _random_cache(shape, dtype: torch.dtype, as_fp8: bool, bound: float, device) -> torch.Tensor:
    """uniform(-bound, bound) in a paged layout.  fp8 caches are drawn in fp16 and narrowed by the cache's own
    converter, so no byte pattern is Inf/NaN (the reference fixture's trick, vllm/utils.py:311-329)."""
    if as_fp8:
        from ._C import cache_ops
        out = torch.empty(shape, dtype=torch.uint8, device=device)
        cache_ops.convert_fp8_e5m2(torch.empty(shape, dtype=torch.float16, device=device).uniform_(-bound, bound), out)
        return out
    if not dtype.is_floating_point:
        raise ValueError(f"Does not support cache of type {dtype}")
    return torch.empty(shape, dtype=dtype, device=device).uniform_(-bound, bound)


def create_kv_caches_with_random(num_blocks: int, block_size: int, num_layers: int, num_heads: int, head_size: int,
                                 cache_dtype: Optional[Union[str, torch.dtype]],
                                 model_dtype: Optional[Union[str, torch.dtype]] = None, seed: Optional[int] = 0,
                                 device: Optional[str] = "cuda") -> Tuple[List[torch.Tensor], List[torch.Tensor]]:
    """Random per-layer (key_caches, value_caches) for tests and benchmarks: same arguments, value range
    (+-head_size^-0.5) and draw order (every key cache, then every value cache) as the reference's fixture of the
    same name (vllm/utils.py:332-397; tests/kernels/conftest.py hands it out as `kv_cache_factory`)."""
    torch.random.manual_seed(seed)
    if torch.cuda.is_available():
        torch.cuda.manual_seed(seed)
    dtype = resolve_cache_dtype(cache_dtype, model_dtype)
    as_fp8 = cache_dtype == "fp8_e5m2"
    bound = head_size ** -0.5
    kshape, vshape = kv_cache_shapes(num_blocks, block_size, num_heads, head_size, dtype)
    keys = [_random_cache(kshape, dtype, as_fp8, bound, device) for _ in range(num_layers)]
    values = [_random_cache(vshape, dtype, as_fp8, bound, device) for _ in range(num_layers)]
    return keys, values


def rotary_cos_sin_cache(rot_dim: int, max_position: int, base: float, dtype: torch.dtype, device="cpu") -> torch.Tensor:
    """[max_position, rot_dim]: rot_dim/2 cosines then rot_dim/2 sines of position * base^(-2i/rot_dim) -- the
    table `rotary_embedding` indexes by position (reference RotaryEmbedding._compute_cos_sin_cache,
    vllm/model_executor/layers/rotary_embedding.py:67-93, cast to the model dtype as in :63-65)."""
    inv_freq = base ** (-torch.arange(0, rot_dim, 2, dtype=torch.float, device=device) / rot_dim)
    angles = torch.outer(torch.arange(max_position, dtype=torch.float, device=device), inv_freq)
    return torch.cat((angles.cos(), angles.sin()), dim=-1).to(dtype)
