"""GPU parity tests of rotary_embedding / batched_rotary_embedding and of the fused
rotary_embedding_reshape_and_cache op, through both bindings of the C ABI.

Checked against (1) the committed outputs of the reference's RotaryEmbedding._forward (tests/golden/rope_ref),
(2) the oracle on seeded inputs in the layouts the model uses (strided q/k/v views of one qkv buffer,
reference tests/kernels/test_pos_encoding.py recipe), (3) the UNMODIFIED reference CUDA kernel (oracle/_ref)
when built.  Bar: bit-exact for f16/bf16 (every product and sum is rounded to the element type on all
sides); fp32 within 1e-6 relative (the fused multiply-add contraction the compiler picks is not part of the
spec) and bit-exact against the reference kernel compiled by the same compiler."""
import pytest
import torch

from _util import bits, load_ref_cuda, load_rope_golden

pytestmark = pytest.mark.gpu

ROPE = load_rope_golden()
DTYPES = [torch.half, torch.bfloat16, torch.float]


def _bindings():
    from vllm_splitwise_b200 import _C
    from ctypes_binding import cache_ops as _ctypes_cache_ops, ops as _ctypes_ops
    out = [("ctypes", _ctypes_ops, _ctypes_cache_ops)]
    if _C.BINDING == "torch":
        out.append(("torch", _C.ops, _C.cache_ops))
    return out


def _same(a, b, dtype):
    if dtype == torch.float:
        return torch.allclose(a, b, rtol=1e-6, atol=1e-7)
    return torch.equal(bits(a), bits(b))


@pytest.mark.parametrize("case", ROPE, ids=[c["meta"]["name"] for c in ROPE])
def test_rotary_golden(case):
    m = case["meta"]
    for name, ops, _ in _bindings():
        q, k = case["q"].cuda(), case["k"].cuda()
        ops.rotary_embedding(case["pos"].cuda(), q, k, m["head_size"], case["cache"].cuda(), m["is_neox"])
        assert _same(q.cpu(), case["oq"], q.dtype) and _same(k.cpu(), case["ok"], k.dtype), name


def _qkv(n, hq, hkv, d, dtype, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    qkv = torch.randn(n, (hq + 2 * hkv) * d, dtype=dtype, device="cuda", generator=g)
    return qkv, qkv[:, :hq * d], qkv[:, hq * d:(hq + hkv) * d], qkv[:, (hq + hkv) * d:]


@pytest.mark.parametrize("is_neox", [True, False])
@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("head_size,rot_dim", [(64, 64), (80, 80), (128, 128), (256, 64), (96, 48)])
@pytest.mark.parametrize("num_tokens", [1, 7, 300])
def test_rotary_vs_oracle_and_reference_kernel(num_tokens, head_size, rot_dim, dtype, is_neox):
    from oracle import paged_kv_oracle as O
    hq, hkv = 8, 2
    cache = O.compute_cos_sin_cache(rot_dim, 2048, 10000, dtype).cuda()
    pos = torch.randint(0, 2048, (num_tokens,), device="cuda", generator=torch.Generator(device="cuda").manual_seed(1))
    ref = load_ref_cuda()
    for name, ops, _ in _bindings():
        qkv, q, k, v = _qkv(num_tokens, hq, hkv, head_size, dtype, 5)
        qkv_o = qkv.cpu().clone()
        v_before = v.clone()
        ops.rotary_embedding(pos, q, k, head_size, cache, is_neox)
        qo, ko = qkv_o[:, :hq * head_size], qkv_o[:, hq * head_size:(hq + hkv) * head_size]
        O.rotary_embedding(pos.cpu(), qo, ko, head_size, cache.cpu(), is_neox)
        assert _same(qkv.cpu(), qkv_o, dtype), name
        assert torch.equal(v, v_before)
        if ref is not None:
            qkv_r, qr, kr, _ = _qkv(num_tokens, hq, hkv, head_size, dtype, 5)
            ref.ops.rotary_embedding(pos, qr, kr, head_size, cache, is_neox)
            assert torch.equal(bits(qkv), bits(qkv_r)), f"{name}: differs from the reference CUDA kernel"


@pytest.mark.parametrize("dtype", [torch.half, torch.bfloat16])
def test_batched_rotary_offsets(dtype):
    from oracle import paged_kv_oracle as O
    n, hq, hkv, d = 33, 4, 4, 128
    cache = O.compute_cos_sin_cache(d, 4096, 10000, dtype).cuda()
    pos = torch.randint(0, 1024, (n,), device="cuda")
    off = torch.randint(0, 3, (n,), device="cuda") * 1024
    ref = load_ref_cuda()
    for name, ops, _ in _bindings():
        qkv, q, k, _ = _qkv(n, hq, hkv, d, dtype, 9)
        qkv2, q2, k2, _ = _qkv(n, hq, hkv, d, dtype, 9)
        ops.batched_rotary_embedding(pos, q, k, d, cache, True, d, off)
        ops.rotary_embedding(pos + off, q2, k2, d, cache, True)
        assert torch.equal(bits(qkv), bits(qkv2)), name
        if ref is not None:
            qkv3, q3, k3, _ = _qkv(n, hq, hkv, d, dtype, 9)
            ref.ops.batched_rotary_embedding(pos, q3, k3, d, cache, True, d, off)
            assert torch.equal(bits(qkv), bits(qkv3)), name


@pytest.mark.parametrize("kv_cache_dtype", ["auto", "fp8_e5m2"])
@pytest.mark.parametrize("is_neox", [True, False])
@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("block_size", [8, 16, 32])
@pytest.mark.parametrize("num_tokens", [1, 40, 200, 600])
@pytest.mark.parametrize("d,rot", [(128, 128), (80, 80), (256, 64)])
def test_fused_equals_the_two_ops(kv_cache_factory, d, rot, num_tokens, block_size, dtype, is_neox, kv_cache_dtype):
    """The fused op IS rotary_embedding followed by reshape_and_cache: same q, k (in place) and caches, bit for
    bit, for decode-sized (single launch) and prefill-sized calls, with padding slots; and the sequence of the
    two reference CUDA kernels gives the same bits when oracle/_ref is built."""
    from oracle import paged_kv_oracle as O
    # (128,128): vector kernel; (80,80) NeoX with an fp8 cache: 40 pairs do not fill 16-element chunks -> scalar
    # kernel; (256,64): pass-through chunks past rot_dim
    hq, hkv, nb = 8, 2, 96
    cache = O.compute_cos_sin_cache(rot, 4096, 10000, dtype).cuda()
    gen = torch.Generator(device="cuda").manual_seed(3)
    pos = torch.randint(0, 4096, (num_tokens,), device="cuda", generator=gen)
    slots = torch.randperm(nb * block_size, device="cuda", generator=gen)[:num_tokens]
    if num_tokens > 2:
        slots[1] = -1
    ref = load_ref_cuda()
    for name, ops, cache_ops in _bindings():
        kcs, vcs = kv_cache_factory(nb, block_size, 1, hkv, d, kv_cache_dtype, dtype, 0, "cuda")
        kc_f, vc_f = kcs[0], vcs[0]
        kc_s, vc_s = kc_f.clone(), vc_f.clone()
        kc_r, vc_r = kc_f.clone(), vc_f.clone()
        qkv_f, qf, kf, vf = _qkv(num_tokens, hq, hkv, d, dtype, 11)
        qkv_s, qs, ks, vs = _qkv(num_tokens, hq, hkv, d, dtype, 11)
        n0 = _launches()
        cache_ops.rotary_embedding_reshape_and_cache(pos, qf, kf, vf, d, cache, is_neox, kc_f, vc_f, slots, kv_cache_dtype)
        fused_launches = _launches() - n0
        ops.rotary_embedding(pos, qs, ks, d, cache, is_neox)
        cache_ops.reshape_and_cache(ks.view(num_tokens, hkv, d), vs.view(num_tokens, hkv, d), kc_s, vc_s, slots,
                                    kv_cache_dtype)
        assert fused_launches == (1 if num_tokens < 512 else 2), (name, fused_launches)
        assert torch.equal(bits(qkv_f), bits(qkv_s)), name
        assert torch.equal(bits(kc_f), bits(kc_s)) and torch.equal(bits(vc_f), bits(vc_s)), name
        if ref is not None:
            qkv_r, qr, kr, vr = _qkv(num_tokens, hq, hkv, d, dtype, 11)
            ref.ops.rotary_embedding(pos, qr, kr, d, cache, is_neox)
            ref.cache_ops.reshape_and_cache(kr.view(num_tokens, hkv, d), vr.view(num_tokens, hkv, d), kc_r, vc_r, slots,
                                            kv_cache_dtype)
            assert torch.equal(bits(qkv_f), bits(qkv_r)), name
            assert torch.equal(bits(kc_f), bits(kc_r)) and torch.equal(bits(vc_f), bits(vc_r)), name
        # the oracle's definition of the fused op on the CPU
        kcs_o, vcs_o = kv_cache_factory(nb, block_size, 1, hkv, d, kv_cache_dtype, dtype, 0, "cuda")
        kc_o, vc_o = kcs_o[0].cpu(), vcs_o[0].cpu()
        qkv_o = _qkv(num_tokens, hq, hkv, d, dtype, 11)[0].cpu()
        O.rotary_embedding_reshape_and_cache(pos.cpu(), qkv_o[:, :hq * d], qkv_o[:, hq * d:(hq + hkv) * d],
                                             qkv_o[:, (hq + hkv) * d:], d, cache.cpu(), is_neox, kc_o, vc_o, slots.cpu(),
                                             kv_cache_dtype)
        if dtype != torch.float:
            assert torch.equal(bits(qkv_f.cpu()), bits(qkv_o)), name
            assert torch.equal(bits(kc_f.cpu()), bits(kc_o)) and torch.equal(bits(vc_f.cpu()), bits(vc_o)), name
        else:
            assert torch.allclose(qkv_f.cpu(), qkv_o, rtol=1e-6, atol=1e-7), name


@pytest.mark.parametrize("dtype", [torch.half, torch.float])
@pytest.mark.parametrize("is_neox", [True, False])
def test_unaligned_rows_take_the_scalar_kernels(dtype, is_neox):
    """q/k/v rows that are not 16-byte aligned (a view one element into a buffer): same results."""
    from oracle import paged_kv_oracle as O
    from vllm_splitwise_b200._C import cache_ops, ops
    n, hq, hkv, d, bs, nb = 9, 4, 2, 64, 16, 16
    cache = O.compute_cos_sin_cache(d, 256, 10000, dtype).cuda()
    pos = torch.arange(n, device="cuda") * 7
    slots = torch.randperm(nb * bs, device="cuda")[:n]
    width = (hq + 2 * hkv) * d
    buf = torch.randn(n * width + 8, dtype=dtype, device="cuda")
    res = []
    for shift in (0, 1):
        qkv = buf[shift:shift + n * width].view(n, width).clone() if shift == 0 else None
        if shift == 1:
            holder = torch.empty(n * width + 8, dtype=dtype, device="cuda")
            qkv = holder[1:1 + n * width].view(n, width)
            qkv.copy_(buf[0:n * width].view(n, width))
        q, k, v = qkv[:, :hq * d], qkv[:, hq * d:(hq + hkv) * d], qkv[:, (hq + hkv) * d:]
        x = 16 // qkv.element_size()
        kc = torch.zeros(nb, hkv, d // x, bs, x, dtype=dtype, device="cuda")
        vc = torch.zeros(nb, hkv, d, bs, dtype=dtype, device="cuda")
        cache_ops.rotary_embedding_reshape_and_cache(pos, q, k, v, d, cache, is_neox, kc, vc, slots, "auto")
        q2 = q.clone()
        ops.rotary_embedding(pos, q2, k.clone(), d, cache, is_neox)   # rotates once more: only exercises the path
        res.append((qkv.clone(), kc, vc, q2))
    for a, b in zip(*res):
        assert torch.equal(bits(a), bits(b))


def _launches():
    from vllm_splitwise_b200 import _lib
    return _lib.load().b200_kv_launch_count()


def test_fused_is_graph_capturable():
    """Decode calls the op inside a CUDA graph (reference model_runner.py:826-834): no allocation, no sync."""
    from oracle import paged_kv_oracle as O
    from vllm_splitwise_b200._C import cache_ops
    n, hq, hkv, d, bs, nb = 16, 8, 2, 128, 16, 32
    cache = O.compute_cos_sin_cache(d, 512, 10000, torch.half).cuda()
    pos = torch.arange(n, device="cuda") * 3
    slots = torch.arange(n, device="cuda") * bs + 5
    qkv, q, k, v = _qkv(n, hq, hkv, d, torch.half, 2)
    qkv0 = qkv.clone()
    kc = torch.zeros(nb, hkv, d // 8, bs, 8, dtype=torch.half, device="cuda")
    vc = torch.zeros(nb, hkv, d, bs, dtype=torch.half, device="cuda")
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        cache_ops.rotary_embedding_reshape_and_cache(pos, q, k, v, d, cache, True, kc, vc, slots, "auto")
    torch.cuda.current_stream().wait_stream(s)
    want_qkv, want_kc, want_vc = qkv.clone(), kc.clone(), vc.clone()
    qkv.copy_(qkv0); kc.zero_(); vc.zero_()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        cache_ops.rotary_embedding_reshape_and_cache(pos, q, k, v, d, cache, True, kc, vc, slots, "auto")
    qkv.copy_(qkv0); kc.zero_(); vc.zero_()
    g.replay()
    torch.cuda.synchronize()
    assert torch.equal(qkv, want_qkv) and torch.equal(kc, want_kc) and torch.equal(vc, want_vc)
