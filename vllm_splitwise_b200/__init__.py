"""B200-native (sm_100a) paged KV-cache hot path of vLLM-splitwise.

Drop-in for the reference's `vllm._C` ops on this path and for its KV-transfer manager:

    from vllm_splitwise_b200._C import ops, cache_ops, cuda_utils      # = vllm._C
    from vllm_splitwise_b200.paged_attn import PagedAttentionImpl      # = ...attention/ops/paged_attn.py
    from vllm_splitwise_b200.cache_engine import CacheEngine           # = vllm/worker/cache_engine.py
    from vllm_splitwise_b200.comm_utils import KVCacheCommManager      # = vllm/worker/comm_utils.py

Everything computes in hand-written CUDA behind the C ABI of `include/b200_kv.h`
(`lib/libb200kv.so`).  There is no CPU or PyTorch fallback: importing `_C` without the built
library raises.
"""
__version__ = "0.1.0"
