"""CPU: host-side mirrors of the reference interface (no device work)."""
import json
import os

import pytest
import torch

from _util import GOLDEN


def test_coalesce_blocks_matches_reference_golden():
    from vllm_splitwise_b200.utils import coalesce_blocks, coalesce_blocks_by_id
    cases = json.load(open(os.path.join(GOLDEN, "coalesce_ref.json")))
    assert len(cases) > 10
    for c in cases:
        assert [list(r) for r in coalesce_blocks(c["blocks"])] == c["runs"]
    d = {i: c["blocks"] for i, c in enumerate(cases[:5])}
    out = coalesce_blocks_by_id(dict(d))
    assert all([list(r) for r in out[i]] == cases[i]["runs"] for i in range(5))
    assert coalesce_blocks([]) == []


def test_v1_v2_selection_rule():
    """reference paged_attn.py:55-63 at the BASELINE configs."""
    from vllm_splitwise_b200.paged_attn import PagedAttentionImpl as P
    assert P.use_v1(4096, 64, 32)        # C2
    assert P.use_v1(4096, 256, 8)        # C3 tp=8
    assert P.use_v1(8192, 128, 32)       # C4
    assert not P.use_v1(8193, 128, 32)   # beyond 8192 -> v2
    assert not P.use_v1(4096, 4, 32)     # small grid, several partitions -> v2
    assert P.use_v1(512, 1, 1)           # one partition -> v1
    assert P.get_supported_head_sizes() == [64, 80, 96, 112, 128, 256]


def test_cache_layout_shapes():
    """reference cache_engine.py:60-75: key [NB,H,D/x,bs,x] with x = 16/elt, value [NB,H,D,bs]."""
    from vllm_splitwise_b200.cache_engine import CacheEngine
    from vllm_splitwise_b200.utils import kv_cache_shapes
    assert kv_cache_shapes(7, 16, 8, 128, torch.half) == ((7, 8, 16, 16, 8), (7, 8, 128, 16))
    assert kv_cache_shapes(7, 32, 8, 128, torch.float) == ((7, 8, 32, 32, 4), (7, 8, 128, 32))
    assert kv_cache_shapes(7, 16, 8, 128, torch.uint8) == ((7, 8, 8, 16, 16), (7, 8, 128, 16))
    # Llama-2-13B tp=4 (C5): 40 layers x (K+V) x 16 x 10 x 128 x 2 B
    assert CacheEngine.get_cache_block_size(16, "auto", torch.half, 128, 10, 40) == 40 * 2 * 40960
    assert CacheEngine.get_cache_block_size(16, "fp8_e5m2", torch.half, 128, 8, 32) == 32 * 2 * 16 * 8 * 128


def test_head_sharding_matches_reference_partition():
    from vllm_splitwise_b200.sharding import shard_heads, stage_ranks
    # C3: Llama-2-70B, 64 q / 8 kv heads
    for tp, (nq, nkv) in {1: (64, 8), 2: (32, 4), 4: (16, 2), 8: (8, 1)}.items():
        shards = [shard_heads(64, 8, tp, r) for r in range(tp)]
        assert all(s.num_q == nq and s.num_kv == nkv for s in shards)
        assert [s.q_start for s in shards] == [r * nq for r in range(tp)]
        assert [s.kv_start for s in shards] == [r * nkv for r in range(tp)]
        # every query head lands on the rank that owns its KV head
        for s in shards:
            for qh in range(s.q_start, s.q_start + s.num_q):
                assert s.kv_start <= qh // 8 < s.kv_start + s.num_kv
    # fewer KV heads than ranks: replication (llama.py:107-111)
    shards = [shard_heads(32, 2, 8, r) for r in range(8)]
    assert [s.kv_start for s in shards] == [0, 0, 0, 0, 1, 1, 1, 1] and all(s.num_kv == 1 for s in shards)
    for s in shards:
        assert all(qh // 16 == s.kv_start for qh in range(s.q_start, s.q_start + s.num_q))
    with pytest.raises(ValueError):
        shard_heads(64, 8, 3, 0)
    assert stage_ranks(8, True) == (range(0, 4), range(4, 8))
    assert stage_ranks(4, False) == (range(4), range(4))


def test_prompt_token_pairing():
    """reference comm_utils.py:165: prompt rank r <-> token rank (r + num_prompt_workers) % world."""
    from vllm_splitwise_b200.comm_utils import is_prompt_rank, pair_rank
    assert [pair_rank(r, 8, 4) for r in range(8)] == [4, 5, 6, 7, 0, 1, 2, 3]
    assert [is_prompt_rank(r, 4) for r in range(8)] == [True] * 4 + [False] * 4
