"""CPU: pins the oracle against the reference's own outputs (tests/golden, produced by
oracle/make_golden.py from /root/reference) and against independent implementations."""
import json
import os

import numpy as np
import pytest
import torch

from oracle import c_oracle
from oracle import paged_kv_oracle as O
from _util import GOLDEN, bits, load_attention_golden

CASES = load_attention_golden()


@pytest.mark.parametrize("case", CASES, ids=[c["meta"]["name"] for c in CASES])
def test_attention_oracle_matches_reference_output(case):
    m = case["meta"]
    out = torch.empty_like(case["q"])
    O.single_query_cached_kv_attention(out, case["q"], m["Hq"] // m["Hkv"], case["kc"], case["vc"], case["bt"],
                                       case["cl"], m["scale"], case["alibi"])
    # same torch ops on identically gathered operands -> expected to be bit-identical; a summation-order
    # change inside a torch build may differ in the last ulp, hence the fallback bound
    if not torch.equal(bits(out), bits(case["out"])):
        tol = 1e-6 if out.dtype == torch.float32 else 2e-3
        assert torch.allclose(out.float(), case["out"].float(), atol=tol, rtol=tol)


@pytest.mark.parametrize("case", [c for c in CASES if c["meta"]["max_ctx"] <= 200],
                         ids=[c["meta"]["name"] for c in CASES if c["meta"]["max_ctx"] <= 200])
def test_loop_and_vectorised_gather_agree(case):
    m = case["meta"]
    # (1) the vectorised gather addresses the cache exactly as the reference's per-token loop does
    bs = case["vc"].shape[3]
    hkv, d = case["vc"].shape[1], case["vc"].shape[2]
    for i in range(m["S"]):
        ctx = int(case["cl"][i])
        k, v = O._gather_sequence(case["kc"], case["vc"], case["bt"][i].tolist(), ctx)
        for j in range(ctx):
            blk, off = int(case["bt"][i][j // bs]), j % bs
            assert torch.equal(bits(k[j]), bits(case["kc"][blk, :, :, off, :].reshape(hkv, d)))
            assert torch.equal(bits(v[j]), bits(case["vc"][blk, :, :, off]))
    # (2) hence the outputs agree (einsum may order its sums differently for differently strided
    # operands, so the last ulp is allowed to move)
    a, b = torch.empty_like(case["q"]), torch.empty_like(case["q"])
    args = (case["q"], m["Hq"] // m["Hkv"], case["kc"], case["vc"], case["bt"], case["cl"], m["scale"], case["alibi"])
    O.single_query_cached_kv_attention(a, *args)
    O.single_query_cached_kv_attention_loop(b, *args)
    tol = 1e-6 if a.dtype == torch.float32 else 2e-3
    assert torch.allclose(a.float(), b.float(), atol=tol, rtol=tol)


@pytest.mark.parametrize("case", CASES, ids=[c["meta"]["name"] for c in CASES])
def test_c_oracle_attention_close_to_reference(case):
    m = case["meta"]
    got = c_oracle.paged_attention(case["q"], case["kc"], case["vc"], case["bt"], case["cl"], m["scale"], case["alibi"])
    f32 = O.attention_fp32(case["q"], m["Hq"] // m["Hkv"], case["kc"], case["vc"], case["bt"], case["cl"], m["scale"],
                           case["alibi"])
    assert torch.allclose(got, f32, atol=2e-6, rtol=1e-5)
    tol = 1e-5 if case["q"].dtype == torch.float32 else 1e-3
    assert torch.allclose(got, case["out"].float(), atol=tol, rtol=1e-3)


def test_v2_partials_recombine_to_full_attention():
    case = [c for c in CASES if c["meta"]["name"] == "gqa_f16_d112_bs32_long"][0]
    m = case["meta"]
    g = m["Hq"] // m["Hkv"]
    P = (m["max_ctx"] + 511) // 512
    es, ml, tmp = O.v2_partials(case["q"], g, case["kc"], case["vc"], case["bt"], case["cl"], m["scale"], None, P)
    full = O.attention_fp32(case["q"], g, case["kc"], case["vc"], case["bt"], case["cl"], m["scale"], None)
    for s in range(m["S"]):
        n = (int(case["cl"][s]) + 511) // 512
        M = ml[s, :, :n].max(dim=1, keepdim=True).values
        w = es[s, :, :n] * torch.exp(ml[s, :, :n] - M)
        rec = (tmp[s, :, :n] * (w / w.sum(dim=1, keepdim=True))[..., None]).sum(dim=1)
        assert torch.allclose(rec, full[s], atol=1e-5, rtol=1e-4)


def test_coalesce_blocks_matches_reference():
    for c in json.load(open(os.path.join(GOLDEN, "coalesce_ref.json"))):
        assert [list(r) for r in O.coalesce_blocks(c["blocks"])] == c["runs"]


def test_e5m2_codec_against_torch_float8_all_fp16_inputs():
    h = np.arange(65536, dtype=np.uint16).view(np.float16)
    f = h.astype(np.float32)
    mine = O.e5m2_from_float32(f)
    theirs = torch.from_numpy(f).to(torch.float8_e5m2).view(torch.uint8).numpy()
    ok = np.isfinite(f) & (np.abs(f) < 61440)        # torch does not saturate: compare below the clamp
    assert np.array_equal(mine[ok], theirs[ok])
    sat = np.isfinite(f) & (np.abs(f) >= 57344)
    assert set((mine[sat] & 0x7F).tolist()) == {0x7B}
    assert set((mine[np.isinf(f)] & 0x7F).tolist()) == {0x7B}
    nn = ~np.isnan(f)
    assert np.array_equal(mine[nn] >> 7, (h.view(np.uint16)[nn] >> 15).astype(np.uint8))  # sign carried unless NaN
    assert set(mine[np.isnan(f)].tolist()) == {0x7F}


def test_e5m2_codec_random_float32_and_c_restatement():
    rs = np.random.RandomState(0)
    f = (rs.randn(1 << 20) * 10.0 ** rs.uniform(-9, 6, 1 << 20)).astype(np.float32)
    mine = O.e5m2_from_float32(f)
    theirs = torch.from_numpy(f).to(torch.float8_e5m2).view(torch.uint8).numpy()
    ok = np.abs(f) < 61440
    assert np.array_equal(mine[ok], theirs[ok])
    assert np.array_equal(mine, c_oracle.e5m2_encode(f))
    allb = np.arange(256, dtype=np.uint8)
    dec_py = O.e5m2_to_float16_bits(allb).view(np.float16).astype(np.float32)
    dec_c = c_oracle.e5m2_decode(allb)
    assert np.array_equal(np.isnan(dec_py), np.isnan(dec_c))
    keep = ~np.isnan(dec_py)
    assert np.array_equal(dec_py[keep], dec_c[keep])
    fin = keep & np.isfinite(dec_py)
    assert np.array_equal(O.e5m2_from_float32(dec_py[fin]), allb[fin])  # decode -> encode is the identity


@pytest.mark.parametrize("dtype", [torch.float32, torch.float16, torch.bfloat16])
@pytest.mark.parametrize("block_size", [8, 16, 32])
def test_reshape_and_cache_python_vs_c(dtype, block_size):
    torch.manual_seed(0)
    n, h, d, nb = 37, 3, 80, 11
    x = 16 // torch.tensor([], dtype=dtype).element_size()
    qkv = torch.randn(n, 3, h, d).to(dtype)
    _, key, value = qkv.unbind(dim=1)  # strided views, as in the reference test (test_cache.py:131-132)
    slots = torch.randperm(nb * block_size)[:n].to(torch.long)
    slots[5] = -1
    kc1 = torch.randn(nb, h, d // x, block_size, x).to(dtype)
    vc1 = torch.randn(nb, h, d, block_size).to(dtype)
    kc2, vc2 = kc1.clone(), vc1.clone()
    O.reshape_and_cache(key, value, kc1, vc1, slots, "auto")
    c_oracle.reshape_and_cache(key, value, kc2, vc2, slots)
    assert torch.equal(bits(kc1), bits(kc2)) and torch.equal(bits(vc1), bits(vc2))
    # round trip: every written token can be read back from its slot
    for i, s in enumerate(slots.tolist()):
        if s < 0:
            continue
        blk, off = s // block_size, s % block_size
        assert torch.equal(kc1[blk, :, :, off, :].reshape(h, d), key[i])
        assert torch.equal(vc1[blk, :, :, off], value[i])


def test_copy_blocks_python_vs_c():
    torch.manual_seed(1)
    kc = torch.randn(20, 2, 8, 16, 8).half()
    ref = kc.clone()
    mapping = {0: [5, 6], 3: [7], 9: [10, 11]}
    O.copy_blocks([ref], [ref.clone()], mapping)
    pairs = [(s, d) for s in sorted(mapping) for d in mapping[s]]
    c_oracle.copy_blocks(kc, pairs)
    assert torch.equal(bits(kc), bits(ref))


# ---------------------------------------------------------------------------------------------
# rotary embedding: the oracle against the reference's RotaryEmbedding._forward
# ---------------------------------------------------------------------------------------------
from _util import load_rope_golden  # noqa: E402

ROPE = load_rope_golden()


@pytest.mark.parametrize("case", ROPE, ids=[c["meta"]["name"] for c in ROPE])
def test_rotary_oracle_matches_reference_forward(case):
    m = case["meta"]
    cache = O.compute_cos_sin_cache(m["rotary_dim"], m["max_position"], m["base"], case["q"].dtype)
    assert torch.equal(bits(cache), bits(case["cache"]))
    q, k = case["q"].clone(), case["k"].clone()
    O.rotary_embedding(case["pos"], q, k, m["head_size"], case["cache"], m["is_neox"])
    # per-operation rounding in the element type on both sides -> bit-identical
    assert torch.equal(bits(q), bits(case["oq"])) and torch.equal(bits(k), bits(case["ok"]))


def test_rotary_oracle_strided_views_and_offsets():
    """In place on the strided q/k views of a fused qkv buffer (how llama.py:155-158 holds them); offsets shift
    the position (batched_rotary_embedding, pos_encoding_kernels.cu:97-123)."""
    torch.manual_seed(0)
    n, hq, hkv, d = 6, 4, 2, 64
    cache = O.compute_cos_sin_cache(d, 64, 10000, torch.float16)
    qkv = torch.randn(n, (hq + 2 * hkv) * d).half()
    dense_q, dense_k = qkv[:, :hq * d].clone(), qkv[:, hq * d:(hq + hkv) * d].clone()
    v_before = qkv[:, (hq + hkv) * d:].clone()
    pos = torch.tensor([0, 5, 17, 3, 9, 30])
    O.rotary_embedding(pos, qkv[:, :hq * d], qkv[:, hq * d:(hq + hkv) * d], d, cache, True)
    O.rotary_embedding(pos, dense_q, dense_k, d, cache, True)
    assert torch.equal(qkv[:, :hq * d], dense_q) and torch.equal(qkv[:, hq * d:(hq + hkv) * d], dense_k)
    assert torch.equal(qkv[:, (hq + hkv) * d:], v_before)
    a, b = dense_q.clone(), dense_k.clone()
    q2, k2 = torch.randn(n, hq * d).half(), torch.randn(n, hkv * d).half()
    q3, k3 = q2.clone(), k2.clone()
    off = torch.tensor([1, 0, 2, 20, 0, 7])
    O.rotary_embedding(pos, q2, k2, d, cache, False, off)
    O.rotary_embedding(pos + off, q3, k3, d, cache, False)
    assert torch.equal(q2, q3) and torch.equal(k2, k3)
