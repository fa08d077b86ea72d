"""TEST INFRASTRUCTURE ONLY.

CPU restatement of the reference's algorithms for the paged-KV hot path.  Only `tests/`,
`__graft_entry__.smoke()` and `bench.py`'s CPU-baseline / `--impl reference` legs may import
anything from this package, and only as the checker or the timed CPU baseline -- never as part of
the product path (`vllm_splitwise_b200/` must not import it; a test enforces that).
"""
