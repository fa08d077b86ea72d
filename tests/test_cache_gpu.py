"""GPU parity tests of the cache ops through the `vllm._C`-shaped module: reshape_and_cache,
copy_blocks, swap_blocks, convert_fp8_e5m2.  Modelled on the reference's tests/kernels/test_cache.py
(:30-226): same grids and input recipes.  Everything here is BIT-EXACT: against the oracle on the
same seeded inputs and against the UNMODIFIED reference CUDA kernels (oracle/_ref) when built.
"""
import random

import numpy as np
import pytest
import torch

from _util import bits, load_ref_cuda

pytestmark = pytest.mark.gpu

DTYPES = [torch.half, torch.bfloat16, torch.float]
HEAD_SIZES = [64, 80, 96, 112, 128, 256]
BLOCK_SIZES = [8, 16, 32]


@pytest.fixture(scope="module")
def cache_ops():
    from vllm_splitwise_b200._C import cache_ops as _c
    return _c


def _same(a, b):
    return torch.equal(bits(a.cpu()), bits(b.cpu()))


def _seed(seed=0):
    random.seed(seed)
    torch.random.manual_seed(seed)
    torch.cuda.manual_seed(seed)


# ---------------------------------------------------------------------------------------------
# reshape_and_cache
# ---------------------------------------------------------------------------------------------
def _rc_inputs(num_tokens, num_heads, head_size, block_size, num_blocks, dtype, kv_cache_dtype, kv_cache_factory,
               pad_every=0, contiguous_slots=False):
    """test_cache.py:118-139: distinct random slots, key/value are strided views of one qkv buffer."""
    num_slots = block_size * num_blocks
    if contiguous_slots:
        start = random.randint(0, num_slots - num_tokens)
        slot_list = list(range(start, start + num_tokens))
    else:
        slot_list = random.sample(range(num_slots), num_tokens)
    if pad_every:
        slot_list = [-1 if i % pad_every == pad_every - 1 else s for i, s in enumerate(slot_list)]
    slot_mapping = torch.tensor(slot_list, dtype=torch.long, device="cuda")
    qkv = torch.randn(num_tokens, 3, num_heads, head_size, dtype=dtype, device="cuda")
    _, key, value = qkv.unbind(dim=1)
    kcs, vcs = kv_cache_factory(num_blocks, block_size, 1, num_heads, head_size, kv_cache_dtype, dtype, 0, "cuda")
    return key, value, kcs[0], vcs[0], slot_mapping


@pytest.mark.parametrize("num_tokens", [42, 1, 700])
@pytest.mark.parametrize("head_size", HEAD_SIZES)
@pytest.mark.parametrize("block_size", BLOCK_SIZES)
@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("kv_cache_dtype", ["auto", "fp8_e5m2"])
def test_reshape_and_cache(cache_ops, kv_cache_factory, num_tokens, head_size, block_size, dtype, kv_cache_dtype):
    from oracle import paged_kv_oracle as O
    _seed()
    key, value, kc, vc, slots = _rc_inputs(num_tokens, 8, head_size, block_size, 1024, dtype, kv_cache_dtype,
                                           kv_cache_factory, pad_every=7 if num_tokens > 1 else 0)
    kc_o, vc_o = kc.cpu().clone(), vc.cpu().clone()
    kc_r, vc_r = kc.clone(), vc.clone()
    cache_ops.reshape_and_cache(key, value, kc, vc, slots, kv_cache_dtype)
    O.reshape_and_cache(key.cpu(), value.cpu(), kc_o, vc_o, slots.cpu(), kv_cache_dtype)
    assert _same(kc, kc_o) and _same(vc, vc_o)
    ref = load_ref_cuda()
    if ref is not None:
        ref.cache_ops.reshape_and_cache(key, value, kc_r, vc_r, slots, kv_cache_dtype)
        assert _same(kc, kc_r) and _same(vc, vc_r)


@pytest.mark.parametrize("kv_cache_dtype", ["auto", "fp8_e5m2"])
@pytest.mark.parametrize("dtype", [torch.half, torch.bfloat16])
def test_reshape_and_cache_prefill_sized(cache_ops, kv_cache_factory, dtype, kv_cache_dtype):
    """Prefill-shaped call (consecutive slots, N = 4096 tokens, not a multiple of the tile) and a fully
    scattered one; both against the reference CUDA kernel (or the oracle on a slice when it is absent)."""
    from oracle import paged_kv_oracle as O
    ref = load_ref_cuda()
    for contiguous in (True, False):
        _seed(1)
        key, value, kc, vc, slots = _rc_inputs(4099, 8, 128, 16, 2048, dtype, kv_cache_dtype, kv_cache_factory,
                                               pad_every=0, contiguous_slots=contiguous)
        kc_r, vc_r = kc.clone(), vc.clone()
        cache_ops.reshape_and_cache(key, value, kc, vc, slots, kv_cache_dtype)
        if ref is not None:
            ref.cache_ops.reshape_and_cache(key, value, kc_r, vc_r, slots, kv_cache_dtype)
        else:
            kc_c, vc_c = kc_r.cpu(), vc_r.cpu()
            O.reshape_and_cache(key.cpu(), value.cpu(), kc_c, vc_c, slots.cpu(), kv_cache_dtype)
            kc_r, vc_r = kc_c, vc_c
        assert _same(kc, kc_r) and _same(vc, vc_r)


@pytest.mark.parametrize("kv_cache_dtype", ["auto", "fp8_e5m2"])
@pytest.mark.parametrize("block_size", [8, 16, 24, 32])
@pytest.mark.parametrize("contiguous", [True, False])
def test_reshape_and_cache_large_prefill_32_token_tiles(cache_ops, kv_cache_factory, block_size, contiguous, kv_cache_dtype):
    """Large prefills (ceil(N / 32) * heads >= 32 * SMs CTAs) take the 32-token tile variant of the kernel
    (csrc/cache_ops.cu launch_reshape_and_cache): 8195 tokens x 32 heads, every block size of the op contract
    plus 24 -- a block size that does not divide the tile, which must fall back to the element-wise V path --
    with consecutive and with scattered slots.  Bit-exact against the reference kernel and the oracle."""
    from oracle import paged_kv_oracle as O
    _seed(5)
    dtype = torch.half
    N, H, D = 8195, 32, 128
    NB = (N + block_size - 1) // block_size + 64
    key, value, kc, vc, slots = _rc_inputs(N, H, D, block_size, NB, dtype, kv_cache_dtype, kv_cache_factory,
                                           pad_every=0 if contiguous else 13, contiguous_slots=contiguous)
    kc_o, vc_o = kc.cpu().clone(), vc.cpu().clone()
    kc_r, vc_r = kc.clone(), vc.clone()
    cache_ops.reshape_and_cache(key, value, kc, vc, slots, kv_cache_dtype)
    O.reshape_and_cache(key.cpu(), value.cpu(), kc_o, vc_o, slots.cpu(), kv_cache_dtype)
    assert _same(kc, kc_o) and _same(vc, vc_o)
    ref = load_ref_cuda()
    if ref is not None:
        ref.cache_ops.reshape_and_cache(key, value, kc_r, vc_r, slots, kv_cache_dtype)
        assert _same(kc, kc_r) and _same(vc, vc_r)


@pytest.mark.parametrize("kv_cache_dtype", ["auto", "fp8_e5m2"])
@pytest.mark.parametrize("dtype,head_size", [(torch.bfloat16, 128), (torch.float, 128), (torch.half, 64), (torch.half, 80)])
@pytest.mark.parametrize("layout", ["block_aligned", "shifted", "shuffled_blocks_padding"])
def test_reshape_and_cache_streamed_prefill(cache_ops, kv_cache_factory, layout, dtype, head_size, kv_cache_dtype):
    """Calls with several (token tile, head) items per resident CTA take the streamed kernel (cp.async ring,
    csrc/cache_ops.cu reshape_and_cache_stream_kernel; 32-token tiles, 16 for fp32): 20007 tokens x 8 heads, a ragged
    last tile, slots that start on a block boundary (vector V path), that start mid-block (element-wise V path), and
    whole blocks in shuffled order with padding tokens in between.  Bit-exact against the oracle and the reference."""
    from oracle import paged_kv_oracle as O
    _seed(9)
    N, H, bs = 20007, 8, 16
    nblk = (N + bs - 1) // bs
    NB = nblk + 9
    key, value, kc, vc, _ = _rc_inputs(N, H, head_size, bs, NB, dtype, kv_cache_dtype, kv_cache_factory,
                                       contiguous_slots=True)
    if layout == "block_aligned":
        slots = torch.arange(N) + 3 * bs
    elif layout == "shifted":
        slots = torch.arange(N) + 3 * bs + 5
    else:
        blocks = torch.randperm(NB)[:nblk]
        slots = (blocks[:, None] * bs + torch.arange(bs)[None, :]).reshape(-1)[:N].clone()
        slots[torch.randperm(N)[:200]] = -1
        slots[64:96] = -1   # a whole tile of padding
    slots = slots.to(torch.long).cuda()
    kc_o, vc_o = kc.cpu().clone(), vc.cpu().clone()
    kc_r, vc_r = kc.clone(), vc.clone()
    cache_ops.reshape_and_cache(key, value, kc, vc, slots, kv_cache_dtype)
    O.reshape_and_cache(key.cpu(), value.cpu(), kc_o, vc_o, slots.cpu(), kv_cache_dtype)
    assert _same(kc, kc_o) and _same(vc, vc_o)
    ref = load_ref_cuda()
    if ref is not None:
        ref.cache_ops.reshape_and_cache(key, value, kc_r, vc_r, slots, kv_cache_dtype)
        assert _same(kc, kc_r) and _same(vc, vc_r)


@pytest.mark.parametrize("kv_cache_dtype", ["auto", "fp8_e5m2"])
@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("block_size", BLOCK_SIZES)
@pytest.mark.parametrize("max_prompt_len", [64, 77])
def test_reshape_and_cache_prompt_batches(cache_ops, kv_cache_factory, max_prompt_len, block_size, dtype, kv_cache_dtype):
    """The prefill slot_mapping the model runner builds (model_runner.py:126-272): [B, max_prompt_len] flattened,
    token i of a prompt at block_table[i // bs] * bs + i % bs, padding -1.  max_prompt_len 64 keeps every tile
    aligned to a block (vector row stores); 77 makes tiles straddle prompts, blocks and padding (element path).
    Also a source whose rows are not 16-byte aligned.  Bit-exact against the oracle and the reference kernel."""
    from oracle import paged_kv_oracle as O
    _seed(3)
    B, H, D, NB = 5, 4, 128, 256
    lens = [max_prompt_len, 1, max_prompt_len - 3, 17, max_prompt_len // 2]
    free = list(range(NB))
    random.shuffle(free)
    slots = []
    for n in lens:
        table = [free.pop() for _ in range((n + block_size - 1) // block_size)]
        row = [table[i // block_size] * block_size + i % block_size for i in range(n)]
        slots += row + [-1] * (max_prompt_len - n)
    slot_mapping = torch.tensor(slots, dtype=torch.long, device="cuda")
    N = B * max_prompt_len
    ref = load_ref_cuda()
    for misalign in (0, 1):
        buf = torch.randn(N * 3 * H * D + 8, dtype=dtype, device="cuda")
        qkv = buf[misalign:misalign + N * 3 * H * D].view(N, 3, H, D)
        _, key, value = qkv.unbind(dim=1)
        kcs, vcs = kv_cache_factory(NB, block_size, 1, H, D, kv_cache_dtype, dtype, 0, "cuda")
        kc, vc = kcs[0], vcs[0]
        kc_o, vc_o = kc.cpu().clone(), vc.cpu().clone()
        kc_r, vc_r = kc.clone(), vc.clone()
        cache_ops.reshape_and_cache(key, value, kc, vc, slot_mapping, kv_cache_dtype)
        O.reshape_and_cache(key.cpu(), value.cpu(), kc_o, vc_o, slot_mapping.cpu(), kv_cache_dtype)
        assert _same(kc, kc_o) and _same(vc, vc_o)
        if ref is not None:
            ref.cache_ops.reshape_and_cache(key, value, kc_r, vc_r, slot_mapping, kv_cache_dtype)
            assert _same(kc, kc_r) and _same(vc, vc_r)


def test_reshape_and_cache_fp8_special_values(cache_ops):
    """Saturation, signed zeros, subnormals, Inf, NaN through the fused fp8 write (quant_utils.cuh:164-192)."""
    from oracle import paged_kv_oracle as O
    vals = torch.tensor([0.0, -0.0, 1.0, -1.0, 57344.0, 60000.0, 65504.0, float("inf"), -float("inf"), float("nan"),
                         6.1e-5, 3.0e-5, 1.5e-5, 7.6e-6, 7.7e-6, 2.3e-5, 1.25, 1.375, 1.5, 1.625, -1.125, 0.1, 1e-8,
                         2.0 ** -17, 2.0 ** -16, 3 * 2.0 ** -17, 49152.0, 53248.0, -61440.0, 61439.0, 0.3333, 448.0],
                        dtype=torch.float32)
    for dtype in DTYPES:
        src = vals.to(dtype).repeat(4).view(1, 1, 128).cuda()
        kc = torch.zeros(2, 1, 8, 16, 16, dtype=torch.uint8, device="cuda")
        vc = torch.zeros(2, 1, 128, 16, dtype=torch.uint8, device="cuda")
        slot = torch.tensor([19], dtype=torch.long, device="cuda")
        cache_ops.reshape_and_cache(src, src, kc, vc, slot, "fp8_e5m2")
        want = O.e5m2_from_tensor(src.cpu()).view(128)
        got_v = vc[1, 0, :, 3].cpu()
        got_k = kc[1, 0, :, 3, :].reshape(128).cpu()
        assert torch.equal(got_v, want), (dtype, got_v, want)
        assert torch.equal(got_k, want)


# ---------------------------------------------------------------------------------------------
# copy_blocks / swap_blocks
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("num_layers", [1, 5, 40])
@pytest.mark.parametrize("head_size", [64, 80, 128])
@pytest.mark.parametrize("block_size", BLOCK_SIZES)
@pytest.mark.parametrize("dtype", [torch.half, torch.float, torch.uint8])
def test_copy_blocks(cache_ops, num_layers, head_size, block_size, dtype):
    from oracle import paged_kv_oracle as O
    _seed()
    num_blocks, num_mappings, num_heads = 1024, 256, 8
    src_blocks = random.sample(range(num_blocks), num_mappings)
    remaining = list(set(range(num_blocks)) - set(src_blocks))
    dst_blocks = random.sample(remaining, 2 * num_mappings)
    mapping = {src_blocks[i]: [dst_blocks[2 * i], dst_blocks[2 * i + 1]] for i in range(num_mappings)}
    x = 16 // torch.tensor([], dtype=dtype).element_size()
    shape_k, shape_v = (num_blocks, num_heads, head_size // x, block_size, x), (num_blocks, num_heads, head_size, block_size)

    def rnd(shape):
        return torch.randint(0, 256, (int(np.prod(shape)) * (16 // x),), dtype=torch.uint8, device="cuda").view(dtype).view(shape)

    kcs, vcs = [rnd(shape_k) for _ in range(num_layers)], [rnd(shape_v) for _ in range(num_layers)]
    kcs_o, vcs_o = [t.clone() for t in kcs], [t.clone() for t in vcs]
    cache_ops.copy_blocks(kcs, vcs, mapping)
    O.copy_blocks(kcs_o, vcs_o, mapping)  # torch ops on the device tensors: the checker
    for a, b in zip(kcs + vcs, kcs_o + vcs_o):
        assert torch.equal(a.view(torch.uint8), b.view(torch.uint8))


def test_copy_blocks_vs_reference_cuda(cache_ops, kv_cache_factory):
    ref = load_ref_cuda()
    if ref is None:
        pytest.skip("oracle/_ref/vllm_ref_C.so not built")
    _seed(2)
    kcs, vcs = kv_cache_factory(512, 16, 4, 8, 128, "auto", torch.bfloat16, 0, "cuda")
    kcs_r, vcs_r = [t.clone() for t in kcs], [t.clone() for t in vcs]
    srcs = random.sample(range(256), 64)
    mapping = {s: [256 + 2 * i, 257 + 2 * i] for i, s in enumerate(srcs)}
    cache_ops.copy_blocks(kcs, vcs, mapping)
    ref.cache_ops.copy_blocks(kcs_r, vcs_r, mapping)
    for a, b in zip(kcs + vcs, kcs_r + vcs_r):
        assert _same(a, b)


@pytest.mark.parametrize("direction", [("cuda", "cpu"), ("cuda", "cuda"), ("cpu", "cuda")])
@pytest.mark.parametrize("head_size", [64, 80, 128, 256])
@pytest.mark.parametrize("block_size", BLOCK_SIZES)
@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("pinned", [True, False])
def test_swap_blocks(cache_ops, kv_cache_factory, direction, head_size, block_size, dtype, pinned):
    """test_cache.py:164-226: distinct src/dst block ids, both K- and V-shaped tensors."""
    _seed()
    src_dev, dst_dev = direction
    num_blocks, num_mappings = 512, 128
    # sources from the lower half of the ids, destinations from the upper half (disjoint, so the
    # in-place cuda->cuda case is well defined); one run of 24 consecutive (src, dst) pairs, which
    # the host<->device path merges into a single copy
    run = {10 + i: 300 + i for i in range(24)}
    srcs = random.sample(sorted(set(range(256)) - set(run)), num_mappings - 24)
    dsts = random.sample(sorted(set(range(256, 512)) - set(run.values())), num_mappings - 24)
    mapping = dict(run)
    mapping.update(zip(srcs, dsts))
    sk, sv = kv_cache_factory(num_blocks, block_size, 1, 8, head_size, "auto", dtype, 0, src_dev)
    dk, dv = kv_cache_factory(num_blocks, block_size, 1, 8, head_size, "auto", dtype, 1, dst_dev)
    if pinned:
        if src_dev == "cpu":
            sk, sv = [sk[0].pin_memory()], [sv[0].pin_memory()]
        if dst_dev == "cpu":
            dk, dv = [dk[0].pin_memory()], [dv[0].pin_memory()]
    if src_dev == dst_dev:  # in place, as the block manager would (one tensor)
        dk, dv = sk, sv
    sk_c, sv_c = sk[0].clone(), sv[0].clone()
    dk_c, dv_c = dk[0].clone(), dv[0].clone()
    cache_ops.swap_blocks(sk[0], dk[0], mapping)
    cache_ops.swap_blocks(sv[0], dv[0], mapping)
    torch.cuda.synchronize()
    for s, d in mapping.items():
        assert _same(sk_c[s], dk[0][d]) and _same(sv_c[s], dv[0][d])
    untouched = sorted(set(range(num_blocks)) - set(mapping.values()))
    idx = torch.tensor(untouched)
    assert _same(dk_c.cpu()[idx], dk[0].cpu()[idx]) and _same(dv_c.cpu()[idx], dv[0].cpu()[idx])


def test_swap_blocks_other_gpu_rejected(cache_ops):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    a = torch.zeros(4, 8, device="cuda:0")
    b = torch.zeros(4, 8, device="cuda:1")
    with pytest.raises(RuntimeError, match="same GPU"):
        cache_ops.swap_blocks(a, b, {0: 1})


# ---------------------------------------------------------------------------------------------
# convert_fp8_e5m2
# ---------------------------------------------------------------------------------------------
def test_convert_fp8_all_fp16_and_bf16_bit_patterns(cache_ops):
    from oracle import paged_kv_oracle as O
    ref = load_ref_cuda()
    allbits = torch.arange(65536, dtype=torch.int32).to(torch.int16)
    for dtype in (torch.half, torch.bfloat16):
        src = allbits.view(dtype).cuda()
        dst = torch.empty(65536, dtype=torch.uint8, device="cuda")
        cache_ops.convert_fp8_e5m2(src, dst)
        want = O.e5m2_from_tensor(src.cpu())
        assert torch.equal(dst.cpu(), want), dtype
        if ref is not None:
            dst_r = torch.empty_like(dst)
            ref.cache_ops.convert_fp8_e5m2(src, dst_r)
            assert torch.equal(dst, dst_r), dtype
    # fp32: random magnitudes across the whole exponent range
    rs = np.random.RandomState(0)
    f = torch.from_numpy((rs.randn(1 << 20) * 10.0 ** rs.uniform(-9, 6, 1 << 20)).astype(np.float32)).cuda()
    dst = torch.empty(1 << 20, dtype=torch.uint8, device="cuda")
    cache_ops.convert_fp8_e5m2(f, dst)
    assert torch.equal(dst.cpu(), O.e5m2_from_tensor(f.cpu()))
    if ref is not None:
        dst_r = torch.empty_like(dst)
        ref.cache_ops.convert_fp8_e5m2(f, dst_r)
        assert torch.equal(dst, dst_r)
    # widening: all 256 codes
    codes = torch.arange(256, dtype=torch.int32).to(torch.uint8).cuda()
    for dtype in DTYPES:
        out = torch.empty(256, dtype=dtype, device="cuda")
        cache_ops.convert_fp8_e5m2(codes, out)
        want = O.e5m2_to_tensor(codes.cpu(), dtype)
        nan = torch.isnan(want)
        assert torch.equal(torch.isnan(out.cpu()), nan)
        assert torch.equal(bits(out.cpu())[~nan], bits(want)[~nan])
        if ref is not None:
            out_r = torch.empty_like(out)
            ref.cache_ops.convert_fp8_e5m2(codes, out_r)
            assert torch.equal(bits(out.cpu())[~nan], bits(out_r.cpu())[~nan])
