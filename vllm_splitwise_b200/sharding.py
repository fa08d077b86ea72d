"""KV-head sharding of the decode path across GPUs: the head partition tensor parallelism already
makes in the reference (vllm/config.py:288-296, vllm/model_executor/models/llama.py:99-111) and the
stage split of Splitwise (vllm/model_executor/parallel_utils/parallel_state.py:92-103).

Attention decode, reshape_and_cache, copy_blocks and swap_blocks are independent per (sequence, KV head):
rank r of tp owns a contiguous range of query heads and the KV heads they map to, its own cache holding
only those heads, and a replica of block_tables / context_lens / slot_mapping
(vllm/worker/model_runner.py:525-547).  There is no collective inside the path.
"""
from typing import NamedTuple, Tuple


class HeadShard(NamedTuple):
    q_start: int      # first global query head of this rank
    num_q: int        # query heads on this rank
    kv_start: int     # first global KV head of this rank
    num_kv: int       # KV heads on this rank (>= 1; replicated when total_kv < tp)


def shard_heads(total_q_heads: int, total_kv_heads: int, tp_size: int, rank: int) -> HeadShard:
    if not 0 <= rank < tp_size:
        raise ValueError(f"rank {rank} outside [0, {tp_size})")
    if total_q_heads % tp_size != 0:
        raise ValueError(f"{total_q_heads} query heads do not divide over tp={tp_size}")
    num_q = total_q_heads // tp_size
    if total_kv_heads >= tp_size:
        if total_kv_heads % tp_size != 0:
            raise ValueError(f"{total_kv_heads} KV heads do not divide over tp={tp_size}")
        num_kv = total_kv_heads // tp_size
        kv_start = rank * num_kv
    else:
        # fewer KV heads than ranks: each KV head is replicated over tp/total_kv ranks
        if tp_size % total_kv_heads != 0:
            raise ValueError(f"tp={tp_size} is not a multiple of {total_kv_heads} KV heads")
        num_kv = 1
        kv_start = rank // (tp_size // total_kv_heads)
    return HeadShard(rank * num_q, num_q, kv_start, num_kv)


def stage_ranks(world_size: int, sep_prompt_token: bool) -> Tuple[range, range]:
    """(prompt ranks, token ranks): Splitwise halves the world (parallel_state.py:92-103,
    vllm/config.py:434-438); without it every rank does both stages."""
    if not sep_prompt_token:
        return range(world_size), range(world_size)
    if world_size % 2 != 0:
        raise ValueError("Splitwise needs an even world size")
    return range(world_size // 2), range(world_size // 2, world_size)
