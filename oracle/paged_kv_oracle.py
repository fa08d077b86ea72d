"""TEST INFRASTRUCTURE ONLY -- the parity oracle (CPU, torch + numpy).

Each function restates, in this repository's own words, the reference algorithm it cites
(file:line relative to the reference tree).  Pinning (see DESIGN.md "Oracle"):
  * attention: checked against the reference's own torch functions
    (tests/kernels/test_attention.py:43-111), AST-extracted from /root/reference by
    oracle/make_golden.py; their outputs are committed under tests/golden/ and re-checked by
    tests/test_oracle_golden.py on every CPU run.
  * coalesce_blocks: same, against vllm/utils.py:403-415.
  * fp8 e5m2 codec: checked on CPU against torch.float8_e5m2 casts for all 65 536 fp16 inputs
    (finite, in range) and on the GPU box bit-for-bit against the reference's compiled
    convert_fp8_e5m2 kernel (oracle/_ref).
  * cache writes / block copies: the reference states them only as inline loops inside GPU tests
    (tests/kernels/test_cache.py:85-91,148-158,222-226); restated here and pinned on the GPU box
    bit-for-bit against the reference's compiled kernels (oracle/_ref).
  * prefix prefill: outputs of the reference's own Triton kernels, run with Triton's CPU interpreter
    (oracle/make_golden_prefix.py -> tests/golden/prefix_prefill_ref.npz).
  * decode metadata: outputs of the reference's own ModelRunner._prepare_decode
    (oracle/make_golden_metadata.py -> tests/golden/decode_metadata_ref.json).
  * rotary embedding: the reference's RotaryEmbedding._forward (oracle/make_golden.py -> tests/golden/rope_ref.*).
  * all-reduce: PARITY UNPINNED by a reference run (multi-GPU kernel); restates custom_all_reduce.cuh:139-166.
  * KV transfer: MSCCL++ (third-party, un-vendored, no pinned version:
    docs/source/splitwise/installing_mscclpp.rst:11-18) -- semantics are "same bytes at the same
    block offsets" (csrc/kv_comm_kernels.cu:15); parity for fragmented lists is UNPINNED by any
    reference test, only the all-blocks constant-fill case is (vllm/worker/worker.py:350-386).
"""
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

PARTITION_SIZE = 512  # reference paged_attn.py:12, attention_kernels.cu:791


# --------------------------------------------------------------------------------------------
# attention
# --------------------------------------------------------------------------------------------
def masked_attention(query: torch.Tensor, key: torch.Tensor, value: torch.Tensor, scale: float,
                     bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    """One sequence: query [1,H,D], key/value [T,H,D] -> [1,H,D].

    Follows ref_masked_attention (tests/kernels/test_attention.py:43-55): logits are the
    dtype-native q.k product scaled and widened to fp32, softmax in fp32, probabilities cast to
    the value dtype, then a dtype-native P.V product.
    """
    logits = (torch.einsum("qhd,khd->hqk", query, key) * scale).float()
    if bias is not None:
        logits = logits + bias.float()
    probs = torch.softmax(logits, dim=-1).to(value.dtype)
    return torch.einsum("hqk,khd->qhd", probs, value)


def _gather_sequence(key_cache: torch.Tensor, value_cache: torch.Tensor, table: Sequence[int],
                     context_len: int) -> Tuple[torch.Tensor, torch.Tensor]:
    """Materialise [T,Hkv,D] keys/values of one sequence from the paged layout
    (key [NB,Hkv,D/x,bs,x], value [NB,Hkv,D,bs]; vllm/worker/cache_engine.py:60-75)."""
    nb, hkv, dx, bs, x = key_cache.shape
    n = (context_len + bs - 1) // bs
    idx = torch.as_tensor(list(table[:n]), dtype=torch.long, device=key_cache.device)
    k = key_cache.index_select(0, idx).permute(0, 3, 1, 2, 4).reshape(n * bs, hkv, dx * x)
    v = value_cache.index_select(0, idx).permute(0, 3, 1, 2).reshape(n * bs, hkv, value_cache.shape[2])
    return k[:context_len], v[:context_len]


def single_query_cached_kv_attention(output: torch.Tensor, query: torch.Tensor,
                                     num_queries_per_kv: int, key_cache: torch.Tensor,
                                     value_cache: torch.Tensor, block_tables: torch.Tensor,
                                     context_lens: torch.Tensor, scale: float,
                                     alibi_slopes: Optional[torch.Tensor]) -> None:
    """Decode attention over the paged cache, one sequence at a time, gathers vectorised.

    Restates ref_single_query_cached_kv_attention (tests/kernels/test_attention.py:58-111):
    token j of sequence i lives in block block_tables[i][j // bs] at offset j % bs; GQA repeats
    each KV head over its query group; the ALiBi bias is slope_h * (j - context_len + 1).
    """
    tables = block_tables.cpu().tolist()
    lens = context_lens.cpu().tolist()
    for i in range(query.shape[0]):
        ctx = int(lens[i])
        keys, values = _gather_sequence(key_cache, value_cache, tables[i], ctx)
        if num_queries_per_kv > 1:
            keys = keys.repeat_interleave(num_queries_per_kv, dim=1)
            values = values.repeat_interleave(num_queries_per_kv, dim=1)
        bias = None
        if alibi_slopes is not None:
            pos = (torch.arange(ctx, device=query.device).int() - ctx + 1).float()
            bias = alibi_slopes.view(-1, 1, 1) * pos.view(1, 1, -1)
        out = masked_attention(query[i:i + 1], keys, values, scale, bias)
        output[i].copy_(out.view(query.shape[1], query.shape[2]))


def single_query_cached_kv_attention_loop(output, query, num_queries_per_kv, key_cache,
                                          value_cache, block_tables, context_lens, scale,
                                          alibi_slopes) -> None:
    """Same result with the reference's literal per-token gather (small cases only); used to show
    that the vectorised gather above addresses the cache exactly as the reference does."""
    bs = value_cache.shape[3]
    hkv, d = value_cache.shape[1], value_cache.shape[2]
    tables = block_tables.cpu().tolist()
    lens = context_lens.cpu().tolist()
    for i in range(query.shape[0]):
        ctx = int(lens[i])
        ks, vs = [], []
        for j in range(ctx):
            blk, off = int(tables[i][j // bs]), j % bs
            ks.append(key_cache[blk, :, :, off, :].reshape(hkv, d))
            vs.append(value_cache[blk, :, :, off])
        keys, values = torch.stack(ks), torch.stack(vs)
        if num_queries_per_kv > 1:
            keys = keys.repeat_interleave(num_queries_per_kv, dim=1)
            values = values.repeat_interleave(num_queries_per_kv, dim=1)
        bias = None
        if alibi_slopes is not None:
            pos = (torch.arange(ctx, device=query.device).int() - ctx + 1).float()
            bias = alibi_slopes.view(-1, 1, 1) * pos.view(1, 1, -1)
        output[i].copy_(masked_attention(query[i:i + 1], keys, values, scale, bias)[0])


def attention_fp32(query, num_queries_per_kv, key_cache, value_cache, block_tables, context_lens,
                   scale, alibi_slopes) -> torch.Tensor:
    """All-fp32 evaluation of the same formula (inputs widened first): the 'true' value both the
    reference's torch path and any CUDA kernel approximate; used to judge which one is closer."""
    out = torch.empty(query.shape, dtype=torch.float32, device=query.device)
    single_query_cached_kv_attention(out, query.float(), num_queries_per_kv, key_cache.float(),
                                     value_cache.float(), block_tables, context_lens, scale,
                                     alibi_slopes)
    return out


def v2_partials(query, num_queries_per_kv, key_cache, value_cache, block_tables, context_lens,
                scale, alibi_slopes, max_num_partitions: int):
    """Per-partition state of the reference's paged_attention_v2 in fp32
    (csrc/attention/attention_kernels.cu:456-476 writing :280-290, 404-416): for partition p over
    tokens [512p, 512p+512): max_logits = max_t logit_t, exp_sums = sum_t e^(logit_t - max),
    tmp_out = softmax-weighted V (normalised by exp_sum + 1e-6).  Entries of partitions past the
    context are left NaN (the kernel does not write them)."""
    s, h, d = query.shape
    dev = query.device
    ml = torch.full((s, h, max_num_partitions), float("nan"), device=dev)
    es = torch.full((s, h, max_num_partitions), float("nan"), device=dev)
    tmp = torch.full((s, h, max_num_partitions, d), float("nan"), device=dev)
    tables = block_tables.cpu().tolist()
    lens = context_lens.cpu().tolist()
    qf = query.float()
    for i in range(s):
        ctx = int(lens[i])
        keys, values = _gather_sequence(key_cache.float(), value_cache.float(), tables[i], ctx)
        keys = keys.repeat_interleave(num_queries_per_kv, dim=1)
        values = values.repeat_interleave(num_queries_per_kv, dim=1)
        logits = torch.einsum("hd,khd->hk", qf[i], keys) * scale
        if alibi_slopes is not None:
            logits = logits + alibi_slopes.view(-1, 1) * (torch.arange(ctx, device=dev) - ctx + 1).float().view(1, -1)
        for p in range((ctx + PARTITION_SIZE - 1) // PARTITION_SIZE):
            lo, hi = p * PARTITION_SIZE, min(ctx, (p + 1) * PARTITION_SIZE)
            lg = logits[:, lo:hi]
            m = lg.max(dim=1).values
            e = torch.exp(lg - m[:, None])
            ssum = e.sum(dim=1)
            ml[i, :, p], es[i, :, p] = m, ssum
            tmp[i, :, p] = torch.einsum("hk,khd->hd", e, values[lo:hi]) / (ssum + 1e-6)[:, None]
    return es, ml, tmp


# --------------------------------------------------------------------------------------------
# fp8 e5m2 codec (unscaled)
# --------------------------------------------------------------------------------------------
E5M2_MAX_NORM = 0x7B  # 57344
E5M2_NAN = 0x7F


def e5m2_from_float32(x: np.ndarray) -> np.ndarray:
    """float32 -> e5m2 byte, round-to-nearest-even, saturate-to-finite.

    Restates what `__nv_cvt_float_to_fp8(x, __NV_SATFINITE, __NV_E5M2)` computes -- the intrinsic
    the reference calls in fp8_e5m2_unscaled::vec_conversion (quant_utils.cuh:164-192): values at
    or above the max normal (57344), and +-Inf, clamp to 0x7B; |x| <= 2^-17 flushes to zero;
    results below 2^-14 are subnormals with ulp 2^-16; NaN gives the canonical NaN 0x7F (the
    sm_89+ `cvt.rn.satfinite.e5m2x2.f32` path; the CUDA header's host emulation would keep two
    payload bits instead).  The sign bit is carried through in every case except NaN, which
    always encodes as +NaN 0x7F on the device (measured on B200 against the compiled reference kernel).
    """
    bits = np.ascontiguousarray(x, dtype=np.float32).view(np.uint32).astype(np.int64)
    sign = (bits >> 31) << 7
    a = bits & 0x7FFFFFFF
    exp = (a >> 23) - 127 + 15          # e5m2-biased exponent
    man = a & 0x7FFFFF
    # normal results: keep 2 mantissa bits, 21 dropped
    res_n = (exp << 2) | (man >> 21)
    rem_n = man & 0x1FFFFF
    up_n = (rem_n > 0x100000) | ((rem_n == 0x100000) & ((res_n & 1) == 1))
    res_n = res_n + up_n
    # subnormal results (exp <= 0): shift in the implicit one
    shift = np.clip(1 - exp, 0, 40)
    full = (man | 0x800000)
    drop = 21 + shift
    res_s = full >> drop
    rem_s = full & ((np.int64(1) << drop) - 1)
    half_s = np.int64(1) << (drop - 1)
    up_s = (rem_s > half_s) | ((rem_s == half_s) & ((res_s & 1) == 1))
    res_s = res_s + up_s
    res = np.where(exp >= 1, res_n, res_s)
    res = np.where(res > E5M2_MAX_NORM, E5M2_MAX_NORM, res)          # overflow / Inf -> max finite
    res = np.where(a <= 0x37000000, 0, res)                          # |x| <= 2^-17 -> 0
    res = res | sign
    res = np.where(a > 0x7F800000, E5M2_NAN, res)                    # NaN: canonical, sign dropped
    return res.astype(np.uint8)


def e5m2_from_tensor(t: torch.Tensor) -> torch.Tensor:
    """fp32/fp16/bf16 tensor -> uint8 e5m2 (widening to fp32 is exact, so one rounding step,
    as in the half/bf16 specialisations at quant_utils.cuh:164-184)."""
    return torch.from_numpy(e5m2_from_float32(t.detach().cpu().float().numpy())).view(t.shape)


def e5m2_to_float16_bits(b: np.ndarray) -> np.ndarray:
    """e5m2 byte -> IEEE half bit pattern: e5m2 is the high byte of a half
    (`__nv_cvt_fp8_to_halfraw`, used at quant_utils.cuh:23-29)."""
    return (b.astype(np.uint16) << 8)


def e5m2_to_tensor(b: torch.Tensor, dtype: torch.dtype) -> torch.Tensor:
    """uint8 e5m2 -> dtype via half (-> float -> bf16 for bf16: quant_utils.cuh:72-81).  Pure bit
    placement + exact widenings, done with torch integer ops so it runs on the tensor's device."""
    h = (b.to(torch.int16) << 8).view(torch.float16)
    return h.float().to(dtype)


def convert_fp8_e5m2(src: torch.Tensor, dst: torch.Tensor) -> None:
    """Whole-cache cast either direction (csrc/cache_kernels.cu:272-320)."""
    if dst.dtype == torch.uint8:
        dst.copy_(e5m2_from_tensor(src))
    else:
        dst.copy_(e5m2_to_tensor(src, dst.dtype))


# --------------------------------------------------------------------------------------------
# cache writes and block moves
# --------------------------------------------------------------------------------------------
def reshape_and_cache(key: torch.Tensor, value: torch.Tensor, key_cache: torch.Tensor,
                      value_cache: torch.Tensor, slot_mapping: torch.Tensor,
                      kv_cache_dtype: str = "auto") -> None:
    """Scatter token i's K/V into slot slot_mapping[i] (negative = padding, skipped).

    key_cache[blk, h, d // x, off, d % x] = key[i, h, d]; value_cache[blk, h, d, off] = value[i, h, d]
    with blk = slot // bs, off = slot % bs (csrc/cache_kernels.cu:154-210; the reference test
    states the same as a loop, tests/kernels/test_cache.py:148-158).  fp8: values go through the
    saturating e5m2 encode first.
    """
    n, h, d = key.shape
    bs, x = key_cache.shape[3], key_cache.shape[4]
    k, v = key, value
    if kv_cache_dtype == "fp8_e5m2":
        k, v = e5m2_from_tensor(key), e5m2_from_tensor(value)
    k = k.reshape(n, h, d // x, x)
    for i, slot in enumerate(slot_mapping.cpu().tolist()):
        if slot < 0:
            continue
        blk, off = slot // bs, slot % bs
        key_cache[blk, :, :, off, :] = k[i]
        value_cache[blk, :, :, off] = v[i]


def compute_cos_sin_cache(rotary_dim: int, max_position: int, base: float, dtype: torch.dtype) -> torch.Tensor:
    """[max_position, rotary_dim]: cos then sin of position * base^(-2i/rotary_dim)
    (reference vllm/model_executor/layers/rotary_embedding.py:67-93, cast as in :63-65)."""
    inv_freq = 1.0 / (base ** (torch.arange(0, rotary_dim, 2, dtype=torch.float) / rotary_dim))
    t = torch.arange(max_position, dtype=torch.float)
    freqs = torch.einsum("i,j -> ij", t, inv_freq)
    return torch.cat((freqs.cos(), freqs.sin()), dim=-1).to(dtype)


def rotary_embedding(positions: torch.Tensor, query: torch.Tensor, key: torch.Tensor, head_size: int,
                     cos_sin_cache: torch.Tensor, is_neox: bool,
                     cos_sin_cache_offsets: Optional[torch.Tensor] = None) -> None:
    """In place on query/key [..., heads * head_size], restating the reference kernel
    (csrc/pos_encoding_kernels.cu:9-95): for pair (x, y) of every head, x' = x*cos - y*sin, y' = y*cos + x*sin,
    NeoX pairs (i, i + rot_dim/2), GPT-J pairs (2i, 2i+1); the rotation covers the first rot_dim elements of a
    head.  Each product and sum is a torch op in the tensor dtype -- float arithmetic rounded to the element type
    per operation, which is what the kernel's scalar_t arithmetic does (and what the reference's own torch
    check, RotaryEmbedding._forward, rotary_embedding.py:95-137, computes)."""
    rot_dim = cos_sin_cache.shape[1]
    embed = rot_dim // 2
    pos = positions.reshape(-1)
    if cos_sin_cache_offsets is not None:
        pos = pos + cos_sin_cache_offsets.reshape(-1)
    cs = cos_sin_cache[pos]                       # [N, rot_dim]
    cos, sin = cs[:, None, :embed], cs[:, None, embed:]
    for t in (query, key):
        n = t.numel() // t.shape[-1]
        v = t.reshape(n, -1, head_size)           # a view for the dense / row-strided layouts the ops accept
        rot = v[..., :rot_dim]
        if is_neox:
            x, y = rot[..., :embed].clone(), rot[..., embed:].clone()
            rot[..., :embed] = x * cos - y * sin
            rot[..., embed:] = y * cos + x * sin
        else:
            x, y = rot[..., 0::2].clone(), rot[..., 1::2].clone()
            rot[..., 0::2] = x * cos - y * sin
            rot[..., 1::2] = y * cos + x * sin
        if v.data_ptr() != t.data_ptr():          # reshape had to copy (non-viewable strides): write back
            t.copy_(v.reshape(t.shape))


def rotary_embedding_reshape_and_cache(positions, query, key, value, head_size, cos_sin_cache, is_neox, key_cache,
                                       value_cache, slot_mapping, kv_cache_dtype) -> None:
    """The fused op is DEFINED as the two reference ops in sequence (models/llama.py:156-158 then
    attention backends -> reshape_and_cache)."""
    rotary_embedding(positions, query, key, head_size, cos_sin_cache, is_neox)
    n = key.numel() // key.shape[-1]
    reshape_and_cache(key.reshape(n, -1, head_size), value.reshape(n, -1, head_size), key_cache, value_cache,
                      slot_mapping.reshape(-1), kv_cache_dtype)


def copy_blocks(key_caches: List[torch.Tensor], value_caches: List[torch.Tensor],
                block_mapping: Dict[int, List[int]]) -> None:
    """For every layer: cache[dst] = cache[src] for K and V (csrc/cache_kernels.cu:65-91;
    tests/kernels/test_cache.py:85-91).  Sources are read from the pre-call state only when a
    block is never both a source and a destination, which the block manager guarantees."""
    for src, dsts in block_mapping.items():
        for dst in dsts:
            for kc in key_caches:
                kc[dst].copy_(kc[src])
            for vc in value_caches:
                vc[dst].copy_(vc[src])


def swap_blocks(src: torch.Tensor, dst: torch.Tensor, block_mapping: Dict[int, int]) -> None:
    """dst[d] = src[s] for each mapping entry (csrc/cache_kernels.cu:21-60;
    tests/kernels/test_cache.py:222-226)."""
    for s, d in block_mapping.items():
        dst[d].copy_(src[s])


def coalesce_blocks(block_list: Sequence[int]) -> List[Tuple[int, int]]:
    """Sorted block ids -> (start, length) runs of consecutive ids (vllm/utils.py:403-420)."""
    runs: List[Tuple[int, int]] = []
    for b in sorted(block_list):
        if runs and runs[-1][0] + runs[-1][1] == b:
            runs[-1] = (runs[-1][0], runs[-1][1] + 1)
        else:
            runs.append((b, 1))
    return runs


def kv_put(src_caches: Sequence[torch.Tensor], dst_caches: Sequence[torch.Tensor],
           runs: Sequence[Tuple[int, int]]) -> None:
    """Prefill->decode transfer semantics: for each registered buffer and each run, the bytes of
    blocks [start, start+len) land at the SAME block offsets of the destination buffer
    (put(dst_mem, off, src_mem, off, bytes), csrc/kv_comm_kernels.cu:15;
    vllm/worker/comm_utils.py:132-147)."""
    for s, d in zip(src_caches, dst_caches):
        for start, n in runs:
            d[start:start + n].copy_(s[start:start + n])


# --------------------------------------------------------------------------------------------
# prefix ("context") prefill attention  
# --------------------------------------------------------------------------------------------
def context_attention_fwd(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, o: torch.Tensor,
                          k_cache: torch.Tensor, v_cache: torch.Tensor, b_loc: torch.Tensor,
                          b_start_loc: torch.Tensor, b_seq_len: torch.Tensor, b_ctx_len: torch.Tensor,
                          max_input_len: int, alibi_slopes: Optional[torch.Tensor] = None,
                          compute_dtype: torch.dtype = torch.float32) -> None:
    """Restates the reference Triton kernels _fwd_kernel / _fwd_kernel_alibi
    (vllm/model_executor/layers/attention/ops/prefix_prefill.py:10-199, 392-618) as dense per-sequence
    attention in `compute_dtype`: new token m of sequence b attends the ctx cached tokens (gathered through
    b_loc from the paged layouts, :87-107) and the new tokens n <= m (:145-149); sm_scale = D**-0.5 (:641);
    ALiBi bias slope*(key position - query position) (:503-508, 572-577).  The grid has ceil(max_input_len/BLOCK)
    row blocks of BLOCK = 128 rows (:635,645): rows past the last started block are left untouched.
    Pinned by tests/golden/prefix_prefill_ref.npz: outputs of the reference kernels themselves, executed in
    the build container with Triton's interpreter (oracle/make_golden_prefix.py)."""
    num_heads, D = q.shape[1], q.shape[2]
    num_kv_heads = k.shape[1]
    G = num_heads // num_kv_heads
    bs = v_cache.shape[3]
    scale = 1.0 / (D ** 0.5)
    for b in range(b_seq_len.shape[0]):
        ctx = int(b_ctx_len[b])
        L = min(int(b_seq_len[b]) - ctx, (int(max_input_len) + 127) // 128 * 128)
        if L <= 0:
            continue
        s0 = int(b_start_loc[b])
        nblk = (ctx + bs - 1) // bs
        blocks = b_loc[b, :nblk].long()
        # K cache [NB, Hkv, D/x, bs, x] -> [ctx, Hkv, D];  V cache [NB, Hkv, D, bs] -> [ctx, Hkv, D]
        kc = k_cache[blocks].permute(0, 3, 1, 2, 4).reshape(nblk * bs, num_kv_heads, D)[:ctx]
        vc = v_cache[blocks].permute(0, 3, 1, 2).reshape(nblk * bs, num_kv_heads, D)[:ctx]
        keys = torch.cat([kc, k[s0:s0 + L]], 0).to(compute_dtype)      # [ctx+L, Hkv, D]
        vals = torch.cat([vc, v[s0:s0 + L]], 0).to(compute_dtype)
        qq = q[s0:s0 + L].to(compute_dtype)                           # [L, H, D]
        keys = keys.repeat_interleave(G, dim=1)
        vals = vals.repeat_interleave(G, dim=1)
        logits = torch.einsum("mhd,nhd->hmn", qq, keys) * scale       # [H, L, ctx+L]
        qpos = ctx + torch.arange(L)
        kpos = torch.arange(ctx + L)
        if alibi_slopes is not None:
            bias = (kpos[None, :] - qpos[:, None]).to(compute_dtype)
            logits = logits + alibi_slopes.to(compute_dtype)[:, None, None] * bias[None]
        logits = logits.masked_fill((kpos[None, :] > qpos[:, None])[None], float("-inf"))
        probs = torch.softmax(logits, dim=-1)
        o[s0:s0 + L] = torch.einsum("hmn,nhd->mhd", probs, vals).to(o.dtype)


# --------------------------------------------------------------------------------------------
# decode metadata  
# --------------------------------------------------------------------------------------------
def prepare_decode(seq_lens: Sequence[int], block_tables: Sequence[Sequence[int]], block_size: int,
                   sliding_window: Optional[int] = None, pad_to: Optional[int] = None,
                   out_blocks: Optional[int] = None):
    """Restates ModelRunner._prepare_decode (vllm/worker/model_runner.py:289-377) for one step: `seq_lens[i]` is
    the sequence length INCLUDING the token being fed (seq_data.get_len(), :303), `block_tables[i]` its physical
    blocks.  Returns (positions int64 [B,1], slot_mapping int64 [B,1], context_lens int32 [B], block_tables int32
    [B, out_blocks]) with the reference's padding: entries past the real batch get position 0, slot -1, context 1
    and an empty table (:336-341); tables are padded with 0 (:370-376).
    Pinned by tests/golden/decode_metadata_ref.json (the reference method itself, executed by
    oracle/make_golden_metadata.py)."""
    positions, slots, ctx, tables = [], [], [], []
    for n, bt in zip(seq_lens, block_tables):
        pos = n - 1
        positions.append([pos])
        ctx.append(n if sliding_window is None else min(n, sliding_window))
        slots.append([bt[pos // block_size] * block_size + pos % block_size])
        bt = list(bt)
        if sliding_window is not None:
            bt = bt[-(sliding_window // block_size):]
        tables.append(bt)
    for _ in range((pad_to or 0) - len(positions)):
        positions.append([0]); slots.append([-1]); ctx.append(1); tables.append([])
    width = out_blocks or max(len(t) for t in tables)
    tables = [t + [0] * (width - len(t)) for t in tables]
    return (torch.tensor(positions, dtype=torch.long), torch.tensor(slots, dtype=torch.long),
            torch.tensor(ctx, dtype=torch.int32), torch.tensor(tables, dtype=torch.int32))


# --------------------------------------------------------------------------------------------
# tensor-parallel all-reduce  
# --------------------------------------------------------------------------------------------
def all_reduce_sum(inputs: Sequence[torch.Tensor]) -> torch.Tensor:
    """Restates the arithmetic of the reference's 1-stage custom all-reduce (csrc/custom_all_reduce.cuh:139-166
    packed_reduce + :215-262): every element is widened to fp32, the ranks' values are added in rank order
    0..world-1, and the sum is narrowed once to the element type.  UNPINNED by a run of the reference here (its
    kernel needs several GPUs; the reference's own test, tests/distributed/test_custom_all_reduce.py, compares
    against NCCL with allclose) -- anchored on that call site and on the published algorithm."""
    acc = torch.zeros(inputs[0].shape, dtype=torch.float32)
    for t in inputs:
        acc = acc + t.to(torch.float32)
    return acc.to(inputs[0].dtype)
