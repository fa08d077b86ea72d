"""TEST INFRASTRUCTURE ONLY.  Generates tests/golden/prefix_prefill_ref.npz from the REFERENCE's own
prefix-prefill kernels (vllm/model_executor/layers/attention/ops/prefix_prefill.py), executed in the
build container with Triton's CPU interpreter (TRITON_INTERPRET=1): the file is loaded from
/root/reference by path (nothing is copied), `torch.cuda.get_device_capability` is stubbed to (8, 0)
so that the reference picks its BLOCK = 128 branch (:635), and the kernels run on CPU tensors.
Inputs follow tests/kernels/test_prefix_prefill.py:33-107 at reduced sizes.

    TRITON_INTERPRET=1 python oracle/make_golden_prefix.py
"""
import importlib.util
import os
import random
import sys

os.environ.setdefault("TRITON_INTERPRET", "1")
import numpy as np  # noqa: E402
import torch  # noqa: E402

REF = os.environ.get("REF", "/root/reference")
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden", "prefix_prefill_ref.npz")


def load_reference():
    path = os.path.join(REF, "vllm/model_executor/layers/attention/ops/prefix_prefill.py")
    spec = importlib.util.spec_from_file_location("ref_prefix_prefill", path)
    mod = importlib.util.module_from_spec(spec)
    torch.cuda.get_device_capability = lambda *a, **k: (8, 0)
    spec.loader.exec_module(mod)
    return mod.context_attention_fwd


def make_case(seed, num_heads, num_kv_heads, head_size, block_size, qlens, ctxlens, alibi, scale=1.0):
    """Builds q/k/v, paged caches and metadata the way test_prefix_prefill.py does (:46-107)."""
    random.seed(seed)
    g = torch.Generator().manual_seed(seed)
    B = len(qlens)
    dtype = torch.float16
    seq_lens = [a + b for a, b in zip(qlens, ctxlens)]
    ntok = sum(qlens)
    q = (torch.rand(ntok, num_heads, head_size, generator=g) * 2 - 1).mul(scale).to(dtype)
    kv = (torch.rand(sum(seq_lens), 2, num_kv_heads, head_size, generator=g) * 2 - 1).mul(scale).to(dtype)
    key, value = kv.unbind(dim=1)
    max_blocks = max((c + block_size - 1) // block_size for c in ctxlens) + 1
    cache_size = B * max_blocks + 3
    k_cache = torch.zeros(cache_size, block_size, num_kv_heads, head_size, dtype=dtype)
    v_cache = torch.zeros_like(k_cache)
    k = torch.zeros(ntok, num_kv_heads, head_size, dtype=dtype)
    v = torch.zeros_like(k)
    perm = torch.randperm(cache_size, generator=g)
    block_table = perm[:B * max_blocks].view(B, max_blocks)
    b_start = np.cumsum([0] + qlens[:-1])
    b_seq_start = np.cumsum([0] + seq_lens[:-1])
    for i in range(B):
        s = int(b_seq_start[i]) + ctxlens[i]
        k[b_start[i]:b_start[i] + qlens[i]] = key[s:s + qlens[i]]
        v[b_start[i]:b_start[i] + qlens[i]] = value[s:s + qlens[i]]
        for blk in range((ctxlens[i] + block_size - 1) // block_size):
            n = min(block_size, ctxlens[i] - blk * block_size)
            src = int(b_seq_start[i]) + blk * block_size
            k_cache[block_table[i, blk], :n] = key[src:src + n]
            v_cache[block_table[i, blk], :n] = value[src:src + n]
    k_cache = k_cache.view(-1, block_size, num_kv_heads, head_size // 8, 8).permute(0, 2, 3, 1, 4).contiguous()
    v_cache = v_cache.view(-1, block_size, num_kv_heads, head_size).permute(0, 2, 3, 1).contiguous()
    slopes = None
    if alibi:
        slopes = torch.tensor([2.0 ** (-(i + 1) * 8.0 / num_heads) for i in range(num_heads)], dtype=torch.float32)
    return dict(q=q, k=k, v=v, k_cache=k_cache, v_cache=v_cache, b_loc=block_table,
                b_start_loc=torch.tensor(b_start, dtype=torch.long), b_seq_len=torch.tensor(seq_lens, dtype=torch.long),
                b_ctx_len=torch.tensor(ctxlens, dtype=torch.long), max_input_len=max(qlens), alibi_slopes=slopes)


CASES = [
    # name, seed, H, Hkv, D, bs, qlens, ctxlens, alibi
    ("mha_d64_bs16", 1, 2, 2, 64, 16, [5, 130], [37, 16], False),
    ("gqa_d128_bs32", 2, 4, 2, 128, 32, [140, 17], [70, 129], False),
    ("mqa_d32_bs8", 3, 4, 1, 32, 8, [33], [9], False),
    ("noctx_d64_bs16", 4, 2, 1, 64, 16, [129, 3], [0, 0], False),
    ("alibi_d64_bs16", 5, 4, 2, 64, 16, [20, 131], [45, 8], True),
    ("d16_bs16", 6, 2, 2, 16, 16, [19], [21], False),
]


def main():
    fwd = load_reference()
    out = {}
    for name, seed, H, Hkv, D, bs, ql, cl, alibi in CASES:
        c = make_case(seed, H, Hkv, D, bs, ql, cl, alibi)
        o = torch.zeros_like(c["q"])
        fwd(c["q"], c["k"], c["v"], o, c["k_cache"], c["v_cache"], c["b_loc"], c["b_start_loc"], c["b_seq_len"],
            c["b_ctx_len"], c["max_input_len"], c["alibi_slopes"])
        assert torch.isfinite(o.float()).all(), name
        out[name] = o.view(torch.int16).numpy()  # fp16 bit patterns
        print(name, "ok, |o| max", float(o.float().abs().max()), flush=True)
    np.savez_compressed(OUT, **out)
    print("wrote", OUT)


if __name__ == "__main__":
    sys.exit(main())
