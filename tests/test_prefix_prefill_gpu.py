"""GPU parity: b200_context_attention_fwd (through the reference-named host function
vllm_splitwise_b200.prefix_prefill.context_attention_fwd) against
  (1) outputs of the reference's own Triton kernels (tests/golden/prefix_prefill_ref.npz), and
  (2) the CPU oracle on seeded inputs, following tests/kernels/test_prefix_prefill.py:33-107.
Tolerances: fp16 atol 1e-3 / rtol 1e-3 (north_star's attention bound); bf16 atol 2e-3 / rtol 1e-2 (the
probabilities are rounded to the 8-bit-mantissa element type before P.V, as tl.dot does in the reference)."""
import os
import random

import numpy as np
import pytest
import torch

from oracle import paged_kv_oracle as O
from oracle.make_golden_prefix import CASES, make_case
from _util import GOLDEN

pytestmark = pytest.mark.gpu

TOL = {torch.float16: (1e-3, 1e-3), torch.bfloat16: (2e-3, 1e-2)}


def _fwd():
    from vllm_splitwise_b200.prefix_prefill import context_attention_fwd
    return context_attention_fwd


def _to_dev(c, dtype=None):
    d = {}
    for k, v in c.items():
        if isinstance(v, torch.Tensor):
            v = v.cuda()
            if dtype is not None and v.is_floating_point() and k != "alibi_slopes":
                v = v.to(dtype)
        d[k] = v
    return d


def _run(c):
    o = torch.zeros_like(c["q"])
    _fwd()(c["q"], c["k"], c["v"], o, c["k_cache"], c["v_cache"], c["b_loc"], c["b_start_loc"], c["b_seq_len"],
           c["b_ctx_len"], c["max_input_len"], c["alibi_slopes"])
    torch.cuda.synchronize()
    return o


def _oracle(c):
    cc = {k: (v.cpu() if isinstance(v, torch.Tensor) else v) for k, v in c.items()}
    o = torch.zeros(cc["q"].shape, dtype=torch.float32)
    O.context_attention_fwd(cc["q"], cc["k"], cc["v"], o, cc["k_cache"], cc["v_cache"], cc["b_loc"], cc["b_start_loc"],
                            cc["b_seq_len"], cc["b_ctx_len"], cc["max_input_len"], cc["alibi_slopes"])
    return o


def _check(out, ref, dtype, what=""):
    atol, rtol = TOL[dtype]
    out, ref = out.float().cpu(), ref.float().cpu()
    assert torch.isfinite(out).all(), f"{what}: non-finite output"
    err = (out - ref).abs()
    bad = err > atol + rtol * ref.abs()
    assert not bad.any(), (f"{what}: {int(bad.sum())} of {bad.numel()} elements off, max |err| {err.max().item():.3e} "
                           f"at {np.unravel_index(int(err.argmax()), err.shape)}")


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_matches_reference_triton_golden(case):
    name, seed, H, Hkv, D, bs, ql, cl, alibi = case
    c = _to_dev(make_case(seed, H, Hkv, D, bs, ql, cl, alibi))
    out = _run(c)
    ref = torch.from_numpy(np.load(os.path.join(GOLDEN, "prefix_prefill_ref.npz"))[name]).view(torch.float16)
    # inputs are in [-1, 1] here: two fp16-P kernels with different summation orders
    out, ref = out.float().cpu(), ref.float()
    assert torch.allclose(out, ref, atol=2e-3, rtol=2e-3), f"max |err| {(out - ref).abs().max().item():.3e}"


def _random_case(seed, H, Hkv, D, bs, B, max_q, max_ctx, dtype, alibi=False, min_q=1, min_ctx=0):
    random.seed(seed)
    ql = [random.randint(min_q, max_q) for _ in range(B)]
    cl = [random.randint(min_ctx, max_ctx) for _ in range(B)]
    c = make_case(seed, H, Hkv, D, bs, ql, cl, alibi, scale=1.0)
    return _to_dev(c, dtype), ql, cl


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("D", [16, 32, 64, 128])
@pytest.mark.parametrize("bs", [8, 16, 32])
@pytest.mark.parametrize("G", [1, 4])
def test_vs_oracle_shapes(dtype, D, bs, G):
    c, ql, cl = _random_case(11 + D + bs + G, 4, 4 // G, D, bs, 4, 300, 200, dtype)
    _check(_run(c), _oracle(c), dtype, f"D={D} bs={bs} G={G} q={ql} ctx={cl}")


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("alibi", [False, True])
def test_vs_oracle_edge_lengths(dtype, alibi):
    # single new token, no context, exact tile multiples, one past a tile, block-boundary contexts
    ql = [1, 1, 128, 129, 64, 65, 127, 256, 3]
    cl = [0, 1, 64, 63, 0, 128, 17, 16, 513]
    c = _to_dev(make_case(5, 8, 2, 128, 16, ql, cl, alibi), dtype)
    _check(_run(c), _oracle(c), dtype, f"alibi={alibi}")


def test_stale_cache_slots_do_not_leak():
    # slots past ctx_len in the last cached block and unused blocks hold NaN / Inf bit patterns
    ql, cl = [70, 9], [21, 130]
    c = make_case(9, 4, 2, 128, 16, ql, cl, False)
    kc = c["k_cache"].clone()
    vc = c["v_cache"].clone()
    nan = torch.tensor(float("nan"), dtype=torch.float16)
    used = torch.zeros(kc.shape[0], 16, dtype=torch.bool)
    for i, ctx in enumerate(cl):
        for t in range(ctx):
            used[c["b_loc"][i, t // 16], t % 16] = True
    kc.permute(0, 3, 1, 2, 4)[~used] = nan
    vc.permute(0, 3, 1, 2)[~used] = float("inf")
    ref = _oracle(c)
    c["k_cache"], c["v_cache"] = kc, vc
    _check(_run(_to_dev(c)), ref, torch.float16, "stale slots")


def test_strided_qkv_views_and_untouched_rows():
    # q, k, v are non-contiguous views of one qkv buffer (llama.py:155-156); rows >= max_input_len stay untouched
    H, Hkv, D = 8, 2, 64
    ql, cl = [40, 150], [33, 5]
    c = _to_dev(make_case(21, H, Hkv, D, 16, ql, cl, False))
    n = c["q"].shape[0]
    qkv = torch.zeros(n, (H + 2 * Hkv) * D, dtype=torch.float16, device="cuda")
    qkv[:, :H * D] = c["q"].reshape(n, -1)
    qkv[:, H * D:(H + Hkv) * D] = c["k"].reshape(n, -1)
    qkv[:, (H + Hkv) * D:] = c["v"].reshape(n, -1)
    qv, kv, vv = qkv.split([H * D, Hkv * D, Hkv * D], dim=-1)
    c2 = dict(c, q=qv.view(n, H, D), k=kv.view(n, Hkv, D), v=vv.view(n, Hkv, D))
    assert not c2["q"].is_contiguous()
    ref = _oracle(c)
    _check(_run(c2), ref, torch.float16, "strided views")
    c3 = dict(c2, max_input_len=100)
    o = torch.full_like(c["q"], 7.0)
    _fwd()(c3["q"], c3["k"], c3["v"], o, c3["k_cache"], c3["v_cache"], c3["b_loc"], c3["b_start_loc"], c3["b_seq_len"],
           c3["b_ctx_len"], 100, None)
    torch.cuda.synchronize()
    # the kernel works in 128-row blocks like the reference (BLOCK = 128): rows of started blocks are computed
    assert (o[40 + 128:] == 7.0).all()
    _check(o[:40 + 128], ref[:40 + 128], torch.float16, "max_input_len")


def test_reference_test_recipe():
    """tests/kernels/test_prefix_prefill.py:33-107 (BS 10, query/context lengths up to 1024, block 32, inputs
    uniform(-1e-3, 1e-3), allclose(atol=1e-6, rtol=0)) with 8 heads instead of 64 to bound the CPU oracle."""
    random.seed(0)
    ql = [random.randint(16, 1024) for _ in range(10)]
    cl = [random.randint(16, 1024) for _ in range(10)]
    for Hkv in (8, 1):
        c = _to_dev(make_case(0, 8, Hkv, 128, 32, ql, cl, False, scale=1e-3))
        out, ref = _run(c).float().cpu(), _oracle(c)
        assert torch.allclose(out, ref, atol=1e-6, rtol=0), f"max |err| {(out - ref).abs().max().item():.3e}"


def test_int32_and_int64_metadata_agree_and_errors():
    c = _to_dev(make_case(3, 4, 2, 64, 16, [50, 7], [20, 90], False))
    o1 = _run(c)
    c32 = dict(c, b_loc=c["b_loc"].int(), b_start_loc=c["b_start_loc"].int(), b_seq_len=c["b_seq_len"].int(),
               b_ctx_len=c["b_ctx_len"].int())
    assert torch.equal(_run(c32), o1)
    with pytest.raises(RuntimeError, match="Unsupported data type"):
        _run(_to_dev(c, torch.float32))
    with pytest.raises(RuntimeError, match="CUDA"):
        _run(dict(c, q=c["q"].cpu()))


def test_forward_prefix_mirror():
    from types import SimpleNamespace

    from vllm_splitwise_b200.paged_attn import PagedAttentionImpl
    c = _to_dev(make_case(4, 4, 4, 128, 16, [90, 31], [64, 10], True))
    md = SimpleNamespace(block_tables=c["b_loc"].int(), start_loc=c["b_start_loc"], prompt_lens=c["b_seq_len"],
                         context_lens=c["b_ctx_len"].int(), max_seq_len=c["max_input_len"])
    out = PagedAttentionImpl.forward_prefix(c["q"], c["k"], c["v"], c["k_cache"], c["v_cache"], md, c["alibi_slopes"])
    _check(out, _oracle(c), torch.float16, "forward_prefix")


# ---- the 256-row kernel (context_attention_tc256.cuh): row-block and tile boundaries ----------------------------
_EDGE_Q = [129, 255, 256, 257, 384, 511]
_EDGE_CTX = [0, 1, 63, 64, 127, 128, 129, 1000]


def _edge_case(seed, ql, cl, dtype, heads=(4, 2), bs=16, alibi=False, D=128):
    c = make_case(seed, heads[0], heads[1], D, bs, ql, cl, alibi, scale=1.0)
    return _to_dev(c, dtype)


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("ctx,bs", [(0, 16), (129, 8), (1000, 32), (63, 16)])
def test_256_row_kernel_head_size_64(ctx, bs, dtype):
    """Head size 64 on the 256-row tcgen05 kernel (N = 64 P.V tiles, four k-steps of Q.K^T): the same boundary lengths."""
    c = _edge_case(300 + ctx, _EDGE_Q, [ctx] * len(_EDGE_Q), dtype, bs=bs, D=64)
    _check(_run(c), _oracle(c), dtype, f"D 64 ctx {ctx} bs {bs}")


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("ctx", _EDGE_CTX)
def test_256_row_kernel_block_and_tile_boundaries(ctx, dtype):
    """Sequences whose new tokens end just before / at / after a 128-row block and a 256-row CTA, over cached prefixes
    that end just before / at / after a 64- and a 128-token tile; one batch per prefix length so that max_input_len
    > 128 routes the call to the 256-row kernel (block B absent, partial and full)."""
    c = _edge_case(100 + ctx, _EDGE_Q, [ctx] * len(_EDGE_Q), dtype)
    _check(_run(c), _oracle(c), dtype, f"ctx {ctx}")


@pytest.mark.parametrize("bs", [8, 32])
def test_256_row_kernel_ragged_batch_alibi_block_sizes(bs):
    """Ragged batch (1 .. 700 new tokens, 0 .. 900 cached) with ALiBi, block sizes 8 and 32, GQA 8:1."""
    random.seed(bs)
    ql = [1, 128, 700] + [random.randint(1, 700) for _ in range(4)]
    cl = [900, 0, 333] + [random.randint(0, 900) for _ in range(4)]
    c = _edge_case(7 + bs, ql, cl, torch.float16, heads=(8, 1), bs=bs, alibi=True)
    _check(_run(c), _oracle(c), torch.float16, f"bs {bs}")


@pytest.mark.parametrize("env", [{"B200_CTX_TN": "64"}, {"B200_CTX_TN": "128"}, {"B200_CTX_ATTN": "tc128"},
                                 {"B200_CTX_PAIR_GROUP": "1"}], ids=["tn64", "tn128", "tc128", "pair_group_1"])
def test_other_kernel_variants_in_a_subprocess(env):
    """The kernel choice is read once per process: both tile sizes of the 256-row kernel for both head sizes (the default
    is 64 tokens for head size 128 and 128 for head size 64), the 128-row kernel forced for long inputs, and the plain CTA
    order run the boundary cases in a child process."""
    import subprocess
    import sys
    child = dict(os.environ, **env)
    r = subprocess.run([sys.executable, "-m", "pytest", "-q", "-x", "-m", "gpu", __file__, "-k",
                        "block_and_tile_boundaries or ragged_batch_alibi or head_size_64"], env=child, capture_output=True, text=True,
                       timeout=600, cwd=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-1000:]
