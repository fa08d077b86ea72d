"""Long contexts, the planner's limits, wide-range queries and concurrent streams (parity holes).

* The reference's own bound: its test draws context lengths up to MAX_SEQ_LEN = max_smem/4 - 512
  (tests/kernels/test_attention.py:17,153-154; ~57 k tokens on B200) because v1 keeps the whole row of logits in
  shared memory.  Above 8192 tokens the op wrapper switches to v2 (paged_attn.py:62-63).  Here v1 has no such
  limit (online softmax), so v1, v2 and the wrapper are all checked up to that bound.
* The v1 split planner: the kMaxSplit = 64 cap, and batches too large for the ticket area (no split).
* bf16 queries over an fp8 cache run the f16 tensor-core path: queries far outside f16's range must not overflow
  or flush (the reference widens K to the query type instead, csrc/attention/dtype_fp8_e5m2.cuh).
* Two streams issuing split v1 calls at the same time, and a graph replay overlapping eager calls: scheduler
  counters and reduce tickets must not be shared between launches that execute concurrently.
"""
import random

import pytest
import torch

from _util import dequantized, make_attention_inputs, oracle_attention, run_v1, run_v2

pytestmark = pytest.mark.gpu

ATOL, RTOL = 1e-3, 1e-3
FP8_ATOL, FP8_RTOL = 1e-2, 1e-3


@pytest.fixture(scope="module")
def ops():
    from vllm_splitwise_b200._C import ops as _ops
    return _ops


def _ref_max_seq_len() -> int:
    """MAX_SEQ_LEN of the reference test: get_max_shared_memory_bytes() // FLOAT32_BYTES - 512."""
    from vllm_splitwise_b200.utils import get_max_shared_memory_bytes
    return get_max_shared_memory_bytes() // 4 - 512


def _close(out, ref, atol=ATOL, rtol=RTOL):
    o, r = out.float(), ref.float()
    return torch.allclose(o, r, atol=atol, rtol=rtol), (o - r).abs().max().item()


class _Meta:
    def __init__(self, a):
        self.block_tables, self.context_lens = a["bt"], a["cl"]
        self.max_context_len, self.kv_cache_dtype = a["max_ctx"], a["kv_cache_dtype"]
        self.slot_mapping = None


def _run(ops, a, how):
    if how == "v1":
        return run_v1(ops, a)
    if how == "v2":
        return run_v2(ops, a)[0]
    from vllm_splitwise_b200.paged_attn import PagedAttentionImpl
    return PagedAttentionImpl.forward_decode(a["q"], a["kc"], a["vc"], _Meta(a), a["hkv"], a["scale"], a["alibi"])


@pytest.mark.parametrize("how", ["v1", "v2", "forward_decode"])
@pytest.mark.parametrize("max_len", [8193, 16384, 32768, "ref_max"])
@pytest.mark.parametrize("num_heads", [(40, 40), (64, 8)])
@pytest.mark.parametrize("dtype,kv_cache_dtype", [(torch.half, "auto"), (torch.bfloat16, "auto"),
                                                  (torch.bfloat16, "fp8_e5m2")])
def test_long_context(ops, how, max_len, num_heads, dtype, kv_cache_dtype):
    max_len = _ref_max_seq_len() if max_len == "ref_max" else max_len
    random.seed(0)
    num_seqs = 7
    ctx = [random.randint(1, max_len) for _ in range(num_seqs)]
    ctx[-1] = max_len
    a = make_attention_inputs(num_seqs=num_seqs, num_heads=num_heads, head_size=128, block_size=16, dtype=dtype,
                              kv_cache_dtype=kv_cache_dtype, num_blocks=4321, context_lens=ctx, use_alibi=False)
    if how == "forward_decode":
        from vllm_splitwise_b200.paged_attn import PagedAttentionImpl
        assert not PagedAttentionImpl.use_v1(max_len, num_seqs, num_heads[0])  # > 8192 -> the v2 branch
    out = _run(ops, a, how)
    atol, rtol = (FP8_ATOL, FP8_RTOL) if kv_cache_dtype == "fp8_e5m2" else (ATOL, RTOL)
    ok, err = _close(out, oracle_attention(a), atol, rtol)
    assert ok, f"max err {err}; ctx {ctx}"


@pytest.mark.parametrize("how", ["v1", "v2"])
@pytest.mark.parametrize("head_size,block_size,dtype", [(64, 32, torch.half), (256, 8, torch.bfloat16),
                                                        (80, 16, torch.float), (112, 32, torch.bfloat16)])
def test_reference_max_seq_len_other_shapes(ops, how, head_size, block_size, dtype):
    """The reference bound at the other head / block sizes of its grid, ALiBi on."""
    if how == "v2" and 512 % block_size:
        pytest.skip("v2 needs block_size | 512")
    max_len = _ref_max_seq_len()
    random.seed(1)
    ctx = [random.randint(1, max_len) for _ in range(3)] + [max_len]
    a = make_attention_inputs(num_seqs=4, num_heads=(16, 4), head_size=head_size, block_size=block_size, dtype=dtype,
                              num_blocks=8000, context_lens=ctx, use_alibi=True, seed=1)
    ok, err = _close(_run(ops, a, how), oracle_attention(a))
    assert ok, f"max err {err}; ctx {ctx}"


@pytest.mark.parametrize("num_seqs,num_heads,ctx_len", [(1, (8, 1), 32768), (1, (32, 32), 16384), (2, (64, 8), 9000),
                                                       (3, (8, 8), 40000)])
def test_v1_split_hits_the_partition_cap(ops, num_seqs, num_heads, ctx_len):
    """Few (sequence, kv-head) items and a long context: the planner wants more partitions than kMaxSplit = 64
    (csrc/paged_attention.cu) and must stop there; partials are combined in-kernel by the ticket holder."""
    import ctypes
    from vllm_splitwise_b200 import _lib
    from ctypes_binding.ops import _fill_args
    a = make_attention_inputs(num_seqs=num_seqs, num_heads=num_heads, head_size=128, block_size=16, dtype=torch.half,
                              num_blocks=num_seqs * (ctx_len // 16 + 1), context_lens=[ctx_len] * num_seqs,
                              distinct_blocks=True, seed=5)
    out = run_v1(ops, a)
    args = _fill_args(out, a["q"], a["kc"], a["vc"], a["hkv"], a["scale"], a["bt"], a["cl"], 16, ctx_len, None, "auto")
    need = _lib.load().b200_paged_attention_workspace_bytes(ctypes.byref(args))
    assert 0 < need < (64 << 20)  # the call is split, and the exact scratch it needs is small
    ok, err = _close(out, oracle_attention(a))
    assert ok, f"max err {err}"
    # the reference's fixed 512-token partitions + separate reduce kernel give the same result
    ok, err = _close(run_v2(ops, a)[0], out)
    assert ok, f"v2 vs v1 max err {err}"


def test_more_items_than_tickets_never_splits(ops):
    """num_seqs * tiles > 65 536 (kTicketInts): the planner must not split (and needs no workspace)."""
    import ctypes
    from vllm_splitwise_b200 import _lib
    from ctypes_binding.ops import _fill_args
    S = 70000
    random.seed(3)
    ctx = [random.randint(1, 300) for _ in range(S)]
    ctx[0], ctx[-1] = 300, 257
    a = make_attention_inputs(num_seqs=S, num_heads=(1, 1), head_size=64, block_size=16, dtype=torch.half,
                              num_blocks=2048, context_lens=ctx, seed=3)
    out = run_v1(ops, a)
    args = _fill_args(out, a["q"], a["kc"], a["vc"], 1, a["scale"], a["bt"], a["cl"], 16, 300, None, "auto")
    assert _lib.load().b200_paged_attention_workspace_bytes(ctypes.byref(args)) == 0
    idx = list(range(0, S, 997)) + [S - 1]
    sub = dict(a, q=a["q"][idx], bt=a["bt"][idx], cl=a["cl"][idx])
    ok, err = _close(out[idx], oracle_attention(sub))
    assert ok, f"max err {err}"
    out2 = run_v2(ops, a)[0]
    ok, err = _close(out2[idx], out[idx])
    assert ok, f"v2 vs v1 max err {err}"


@pytest.mark.parametrize("num_heads", [(16, 2), (32, 8), (48, 1)])
@pytest.mark.parametrize("how", ["v1", "v2"])
def test_bf16_query_outside_f16_range_with_fp8_cache(ops, num_heads, how):
    """|q| up to 1e5 (> f16 max 65 504) and down to 1e-8 (< f16's smallest subnormal) per head.  Checked against
    the all-fp32 evaluation of the formula: with logits of magnitude 1e5 the bf16-native einsum of the torch
    reference cannot resolve the arg-max, the fp32 dot product of the reference CUDA kernel
    (attention_utils.cuh:31-37) can."""
    from oracle import paged_kv_oracle as O
    hq, hkv = num_heads
    ctx = [1, 17, 300, 1025, 2500]
    a = make_attention_inputs(num_seqs=len(ctx), num_heads=num_heads, head_size=128, block_size=16,
                              dtype=torch.bfloat16, kv_cache_dtype="fp8_e5m2", num_blocks=1024, context_lens=ctx, seed=11)
    mags = [1e5, 3e4, 1.0, 1e-3, 1e-8, 6.6e4, 7e4, 1e-6, 2.5e38 ** 0.5, 1e-20]
    scale_per_head = torch.tensor([mags[h % len(mags)] for h in range(hq)], dtype=torch.float32, device="cuda")
    a["q"] = (a["q"].float() / a["scale"] * scale_per_head[None, :, None]).to(torch.bfloat16)
    # one head with both extremes in the same row
    a["q"][:, 0, ::2] = (a["q"][:, 0, ::2].float() * 1e-12).to(torch.bfloat16)
    out = _run(ops, a, how)
    assert torch.isfinite(out).all()
    kc, vc = dequantized(a)
    exact = O.attention_fp32(a["q"], hq // hkv, kc, vc, a["bt"], a["cl"], a["scale"], None)
    ok, err = _close(out, exact, 2e-3, 1e-2)
    assert ok, f"max err {err}"


def _split_case(seed, ctx_len=4096, num_heads=(8, 1), num_seqs=2):
    return make_attention_inputs(num_seqs=num_seqs, num_heads=num_heads, head_size=128, block_size=16, dtype=torch.half,
                                 num_blocks=num_seqs * ((ctx_len + 15) // 16), context_lens=[ctx_len] * num_seqs,
                                 distinct_blocks=True, seed=seed)


def test_two_streams_issue_split_v1_calls_concurrently(ops):
    """Both calls are split by the planner (2 items for 296 resident CTAs) and need reduce tickets + partials;
    issued from two streams they execute at the same time.  Every result must match the oracle."""
    cases = [_split_case(21), _split_case(22, ctx_len=6000, num_heads=(16, 2), num_seqs=1)]
    refs = [oracle_attention(a) for a in cases]
    streams = [torch.cuda.Stream(), torch.cuda.Stream()]
    outs = [[torch.empty_like(a["q"]) for _ in range(40)] for a in cases]
    torch.cuda.synchronize()
    for it in range(40):
        for a, s, o in zip(cases, streams, outs):
            with torch.cuda.stream(s):
                run_v1(ops, a, o[it])
    torch.cuda.synchronize()
    for a, ref, o in zip(cases, refs, outs):
        for it in range(40):
            ok, err = _close(o[it], ref)
            assert ok, f"iteration {it}: max err {err}"


def test_graph_replay_overlapping_eager_calls(ops):
    """A captured split call replayed on one stream while eager split calls run on another: the captured launch
    owns its work counter and its workspace (one per capture), so neither disturbs the other."""
    a, b = _split_case(31), _split_case(32, ctx_len=5000, num_heads=(32, 4), num_seqs=1)
    ref_a, ref_b = oracle_attention(a), oracle_attention(b)
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    out_a = torch.empty_like(a["q"])
    with torch.cuda.stream(s1):
        run_v1(ops, a, out_a)  # warm up outside the capture
    s1.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=s1):
        run_v1(ops, a, out_a)
    outs_b = [torch.empty_like(b["q"]) for _ in range(30)]
    snaps_a = []
    torch.cuda.synchronize()
    for it in range(30):
        with torch.cuda.stream(s1):
            g.replay()
            snaps_a.append(out_a.clone())
        with torch.cuda.stream(s2):
            run_v1(ops, b, outs_b[it])
    torch.cuda.synchronize()
    for it in range(30):
        ok, err = _close(snaps_a[it], ref_a)
        assert ok, f"graph replay {it}: max err {err}"
        ok, err = _close(outs_b[it], ref_b)
        assert ok, f"eager call {it}: max err {err}"


def _quad_ragged_case(kv_cache_dtype, block_size, use_alibi, seed):
    """Enough (sequence, head tile, partition) items for the v2 quad mode (>= 1.5 quads per resident CTA), with context
    lengths on and around the 512-token partition edges, odd and even partition counts, and one-token contexts."""
    rng = random.Random(seed)
    edge = [1, 15, 511, 512, 513, 1023, 1024, 1025, 1535, 1536, 1537, 2047, 2048, 2049, 3071, 3072, 3073, 4095, 4096]
    ctx = edge + [rng.randint(1, 4096) for _ in range(256 - len(edge))]
    ctx[-1] = 4096
    return make_attention_inputs(num_seqs=256, num_heads=(16, 2), head_size=128, block_size=block_size,
                                 dtype=torch.bfloat16, kv_cache_dtype=kv_cache_dtype, num_blocks=4096, context_lens=ctx,
                                 use_alibi=use_alibi, seed=seed)


@pytest.mark.parametrize("kv_cache_dtype,block_size,use_alibi", [("auto", 16, False), ("fp8_e5m2", 16, True),
                                                                 ("fp8_e5m2", 32, False)])
def test_v2_quad_mode_ragged_partitions(ops, kv_cache_dtype, block_size, use_alibi):
    """v2 signature on the tensor-core consumer, quad mode (every warp owns whole partitions): output, and the
    partials of every partition a sequence has, against the oracle.  Run again with two partitions per warp-item
    forced by test_v2_pair_mode_in_a_subprocess."""
    from oracle import paged_kv_oracle as O
    a = _quad_ragged_case(kv_cache_dtype, block_size, use_alibi, seed=21)
    out, es, ml, tmp = run_v2(ops, a)
    ref = oracle_attention(a)
    atol, rtol = (FP8_ATOL, FP8_RTOL) if kv_cache_dtype == "fp8_e5m2" else (ATOL, RTOL)
    ok, err = _close(out, ref, atol, rtol)
    assert ok, f"out max err {err}"
    kc, vc = dequantized(a)
    es_o, ml_o, tmp_o = O.v2_partials(a["q"], a["hq"] // a["hkv"], kc, vc, a["bt"], a["cl"], a["scale"], a["alibi"],
                                      es.shape[2])
    cl = a["cl"].tolist()
    for s in range(0, len(cl), 7):
        n = (cl[s] + 511) // 512
        assert torch.allclose(ml[s, :, :n].cpu(), ml_o[s, :, :n].cpu(), atol=2e-3, rtol=1e-3), f"max_logits seq {s}"
        assert torch.allclose(es[s, :, :n].cpu(), es_o[s, :, :n].cpu(), atol=3e-3, rtol=3e-3), f"exp_sums seq {s}"
        ok, err = _close(tmp[s, :, :n].cpu(), tmp_o[s, :, :n].cpu(), atol, rtol)
        assert ok, f"tmp_out seq {s} max err {err}"


def test_v2_pair_mode_in_a_subprocess():
    """Two partitions per warp-item are only used from six doubled quads per resident CTA (full-size C4 in
    test_vs_reference_cuda_full_gpu.py); B200_ATTN_PAIR_MIN_QUADS=0 forces them for every quad-mode call in a child
    process, which re-runs the ragged-partition test above, the guard-region quad case and the v2 tests."""
    import os
    import subprocess
    import sys
    child = dict(os.environ, B200_ATTN_PAIR_MIN_QUADS="0")
    here = os.path.dirname(os.path.abspath(__file__))
    r = subprocess.run([sys.executable, "-m", "pytest", "-q", "-x", "-m", "gpu", os.path.join(here, "test_attention_long_gpu.py"),
                        os.path.join(here, "test_guard_regions_gpu.py"), "-k", "quad_mode_ragged or v2_quad"],
                       env=child, capture_output=True, text=True, timeout=900, cwd=os.path.dirname(here))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-1000:]
