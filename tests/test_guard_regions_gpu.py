"""Out-of-bounds WRITE detection without compute-sanitizer (the tool is closed on this GPU pool: every invocation
exits 86, profiles/r05_sanitizer_unavailable.txt).  Every tensor a kernel writes is carved out of a larger allocation
with a guard region of a known byte pattern on both sides; after the op the guards must be untouched and the result
must still match the oracle / the plain allocation.  Covers the writers of the hot path: reshape_and_cache (token and
tile kernels), paged_attention v1 (direct, split with workspace) and v2 (item and quad mode: out, tmp_out, exp_sums,
max_logits), copy_blocks, convert_fp8_e5m2, the KV put kernels (both variants) and context_attention_fwd.
Out-of-bounds READS are not detectable this way; the block-table / slot / run indices that drive every read are the
ones the parity tests pin bit-for-bit."""
import random

import pytest
import torch

from _util import bits, make_attention_inputs, oracle_attention, run_v1

pytestmark = pytest.mark.gpu

GUARD = 1 << 20  # bytes on each side
PATTERN = 0xA5


class Guarded:
    """A tensor in the middle of a byte buffer whose first and last GUARD bytes hold PATTERN."""

    def __init__(self, shape, dtype, device="cuda", fill=None):
        n = 1
        for s in shape:
            n *= s
        nbytes = n * torch.tensor([], dtype=dtype).element_size()
        self.raw = torch.full((2 * GUARD + nbytes,), PATTERN, dtype=torch.uint8, device=device)
        self.t = self.raw[GUARD:GUARD + nbytes].view(dtype).view(shape)
        if fill is not None:
            self.t.copy_(fill)
        self.nbytes = nbytes

    def intact(self) -> bool:
        return bool((self.raw[:GUARD] == PATTERN).all()) and bool((self.raw[GUARD + self.nbytes:] == PATTERN).all())


def _ok(a, b, atol=1e-3):
    return torch.allclose(a.float(), b.float(), atol=atol, rtol=1e-3)


@pytest.mark.parametrize("num_tokens,contiguous", [(37, False), (700, True), (8200, True), (4099, False),
                                                   (16400, True), (12001, False)])  # the last two: streamed kernel
@pytest.mark.parametrize("kv", ["auto", "fp8_e5m2"])
def test_reshape_and_cache_writes_stay_inside_the_caches(num_tokens, contiguous, kv):
    from oracle import paged_kv_oracle as O
    from vllm_splitwise_b200._C import cache_ops
    torch.manual_seed(0)
    H, D, bs = 8, 128, 16
    NB = (num_tokens + bs - 1) // bs + 3 if contiguous else num_tokens + 3
    cdt = torch.uint8 if kv == "fp8_e5m2" else torch.half
    x = 16 // (1 if kv == "fp8_e5m2" else 2)
    kc = Guarded((NB, H, D // x, bs, x), cdt, fill=torch.zeros(NB, H, D // x, bs, x, dtype=cdt, device="cuda"))
    vc = Guarded((NB, H, D, bs), cdt, fill=torch.zeros(NB, H, D, bs, dtype=cdt, device="cuda"))
    key = torch.empty(num_tokens, H, D, dtype=torch.half, device="cuda").uniform_(-1, 1)
    val = torch.empty_like(key).uniform_(-1, 1)
    if contiguous:   # the LAST block of the cache is used: an overrun would land in the guard
        blocks = torch.randperm(NB - 1)[:(num_tokens + bs - 1) // bs - 1].tolist() + [NB - 1]
        slots = torch.tensor([b * bs + o for b in blocks for o in range(bs)][:num_tokens])
    else:
        slots = torch.randperm(NB)[:num_tokens] * bs + torch.randint(0, bs, (num_tokens,))
        slots[0] = (NB - 1) * bs + bs - 1
        slots[1] = 0
    slots = slots.to(torch.int64).cuda()
    cache_ops.reshape_and_cache(key, val, kc.t, vc.t, slots, kv)
    torch.cuda.synchronize()
    assert kc.intact() and vc.intact()
    kr, vr = torch.zeros_like(kc.t).cpu(), torch.zeros_like(vc.t).cpu()
    O.reshape_and_cache(key.cpu(), val.cpu(), kr, vr, slots.cpu(), kv)
    assert torch.equal(bits(kc.t.cpu()), bits(kr)) and torch.equal(bits(vc.t.cpu()), bits(vr))


@pytest.mark.parametrize("case", ["mha_v1", "gqa_split_v1", "gqa_v2_items", "gqa_fp8_v2_quad"])
def test_attention_writes_stay_inside_their_outputs(case):
    from vllm_splitwise_b200._C import ops
    if case == "mha_v1":
        a = make_attention_inputs(num_seqs=7, num_heads=(8, 8), head_size=128, block_size=16, dtype=torch.half,
                                  num_blocks=512, context_lens=[1, 15, 16, 17, 700, 1023, 2049], seed=1)
    elif case == "gqa_split_v1":   # 2 base items on 296 resident CTAs: the planner splits, partials go to the workspace
        a = make_attention_inputs(num_seqs=2, num_heads=(8, 1), head_size=128, block_size=16, dtype=torch.half,
                                  num_blocks=1024, context_lens=[4096, 3000], seed=2)
    elif case == "gqa_v2_items":
        a = make_attention_inputs(num_seqs=5, num_heads=(16, 2), head_size=128, block_size=16, dtype=torch.bfloat16,
                                  num_blocks=1024, context_lens=[1, 511, 512, 513, 3000], seed=3)
    else:                          # enough partitions for quad mode (>= 1.5 quads per resident CTA)
        rng = random.Random(4)
        ctx = [rng.randint(3000, 4096) for _ in range(96)]
        ctx[0], ctx[-1] = 4096, 3585
        a = make_attention_inputs(num_seqs=96, num_heads=(16, 4), head_size=128, block_size=16, dtype=torch.bfloat16,
                                  kv_cache_dtype="fp8_e5m2", num_blocks=4096, context_lens=ctx, seed=4)
    ref = oracle_attention(a)
    atol = 1e-2 if a["kv_cache_dtype"] == "fp8_e5m2" else 1e-3
    q = a["q"]
    S, Hq, D = q.shape
    out = Guarded((S, Hq, D), q.dtype)
    if case.endswith("v1"):
        run_v1(ops, a, out.t)
        torch.cuda.synchronize()
        assert out.intact()
    else:
        P = (a["max_ctx"] + 511) // 512
        tmp, es, ml = Guarded((S, Hq, P, D), q.dtype), Guarded((S, Hq, P), torch.float32), Guarded((S, Hq, P), torch.float32)
        ops.paged_attention_v2(out.t, es.t, ml.t, tmp.t, q, a["kc"], a["vc"], a["hkv"], a["scale"], a["bt"], a["cl"],
                               a["block_size"], a["max_ctx"], a["alibi"], a["kv_cache_dtype"])
        torch.cuda.synchronize()
        assert out.intact() and tmp.intact() and es.intact() and ml.intact()
    assert _ok(out.t, ref, atol), (out.t.float() - ref.float()).abs().max()


def test_copy_blocks_and_convert_stay_inside():
    from vllm_splitwise_b200._C import cache_ops
    torch.manual_seed(5)
    L, NB, H, D, bs = 3, 40, 4, 64, 16
    ks = [Guarded((NB, H, D // 8, bs, 8), torch.half, fill=torch.randn(NB, H, D // 8, bs, 8, device="cuda").half()) for _ in range(L)]
    vs = [Guarded((NB, H, D, bs), torch.half, fill=torch.randn(NB, H, D, bs, device="cuda").half()) for _ in range(L)]
    before = [g.t.clone() for g in ks + vs]
    mapping = {0: [NB - 1, 7], 5: [6], NB - 2: [1]}
    cache_ops.copy_blocks([g.t for g in ks], [g.t for g in vs], mapping)
    torch.cuda.synchronize()
    assert all(g.intact() for g in ks + vs)
    for g, b in zip(ks + vs, before):
        want = b.clone()
        for s, ds in mapping.items():
            for d in ds:
                want[d] = b[s]
        assert torch.equal(bits(g.t), bits(want))
    n = 16 * 1000 + 5   # a tail that is not a whole 16-element group
    src = torch.randn(n, device="cuda").half()
    dst = Guarded((n,), torch.uint8)
    cache_ops.convert_fp8_e5m2(src, dst.t)
    back = Guarded((n,), torch.half)
    cache_ops.convert_fp8_e5m2(dst.t, back.t)
    torch.cuda.synchronize()
    assert dst.intact() and back.intact()
    assert torch.equal(dst.t, src.to(torch.float8_e5m2).view(torch.uint8))


@pytest.mark.parametrize("variant", [0, 1])
def test_kv_put_writes_only_the_named_blocks(variant):
    import ctypes
    from vllm_splitwise_b200 import _lib
    from vllm_splitwise_b200.comm_utils import KVCacheCommManager
    lib = _lib.load()
    v0, c0 = ctypes.c_int32(), ctypes.c_int32()
    lib.b200_kv_put_config(ctypes.byref(v0), ctypes.byref(c0))
    lib.b200_kv_put_configure(variant, 0)
    try:
        L, NB, H, D, bs = 3, 64, 10, 128, 16
        g = torch.Generator(device="cuda").manual_seed(6)
        src = [(torch.randint(-30000, 30000, (NB, H, D // 8, bs, 8), dtype=torch.int16, device="cuda", generator=g).view(torch.half),
                torch.randint(-30000, 30000, (NB, H, D, bs), dtype=torch.int16, device="cuda", generator=g).view(torch.half))
               for _ in range(L)]
        dk = [Guarded((NB, H, D // 8, bs, 8), torch.half, fill=torch.zeros(NB, H, D // 8, bs, 8, dtype=torch.half, device="cuda")) for _ in range(L)]
        dv = [Guarded((NB, H, D, bs), torch.half, fill=torch.zeros(NB, H, D, bs, dtype=torch.half, device="cuda")) for _ in range(L)]
        m = KVCacheCommManager(rank=0, world_size=2, num_prompt_workers=1)
        m.setup_comm(L, src, peer_kv_cache=[(a.t, b.t) for a, b in zip(dk, dv)], use_comm_stream=True)
        runs = ((0, 1), (10, 5), (NB - 3, 3))       # first block, a run, and the run that ends at the last block
        for layer in range(L):
            m.put_runs(0, layer, runs)
        m.signal_and_flush(0)
        m.join_current_stream()
        m.wait(0)
        torch.cuda.synchronize()
        assert all(x.intact() for x in dk + dv)
        moved = torch.zeros(NB, dtype=torch.bool, device="cuda")
        for s, n in runs:
            moved[s:s + n] = True
        for (sk, sv), a, b in zip(src, dk, dv):
            for s_, d_ in ((sk, a.t), (sv, b.t)):
                assert torch.equal(bits(d_[moved]), bits(s_[moved])) and not bool(d_[~moved].any())
        m.destroy_comm()
    finally:
        lib.b200_kv_put_configure(v0.value, c0.value)


def test_context_attention_writes_stay_inside_the_output():
    from oracle import paged_kv_oracle as O
    from vllm_splitwise_b200.prefix_prefill import context_attention_fwd
    torch.manual_seed(7)
    Hq, Hkv, D, bs, NB = 8, 2, 128, 16, 64
    ql, cl = [130, 9, 257], [21, 130, 0]
    n = sum(ql)
    scale = D ** -0.5
    mk = lambda *s: torch.empty(*s, dtype=torch.half, device="cuda").uniform_(-scale, scale)  # noqa: E731
    q, k, v = mk(n, Hq, D), mk(n, Hkv, D), mk(n, Hkv, D)
    kc, vc = mk(NB, Hkv, D // 8, bs, 8), mk(NB, Hkv, D, bs)
    bt = torch.stack([torch.randperm(NB)[:16] for _ in ql]).int().cuda()
    start = torch.tensor([0, ql[0], ql[0] + ql[1]], dtype=torch.int32, device="cuda")
    seq = torch.tensor([a + b for a, b in zip(ql, cl)], dtype=torch.int32, device="cuda")
    ctx = torch.tensor(cl, dtype=torch.int32, device="cuda")
    out = Guarded((n, Hq, D), torch.half, fill=torch.zeros(n, Hq, D, dtype=torch.half, device="cuda"))
    context_attention_fwd(q, k, v, out.t, kc, vc, bt, start, seq, ctx, max(ql), None)
    torch.cuda.synchronize()
    assert out.intact()
    ref = torch.zeros(n, Hq, D, dtype=torch.float32)
    O.context_attention_fwd(q.cpu(), k.cpu(), v.cpu(), ref, kc.cpu(), vc.cpu(), bt.cpu(), start.cpu(), seq.cpu(), ctx.cpu(),
                            max(ql), None)
    assert _ok(out.t.cpu(), ref)
