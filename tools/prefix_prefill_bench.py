"""Prefix-prefill (context attention) throughput on one GPU: our sm_100a kernel through the reference-named
host function, against torch SDPA on the same sequences with the cached prefix gathered into dense tensors
(a library baseline; the reference's Triton kernel cannot travel to the GPU box).  One JSON line per case.

flops = 4 * D * H * sum_b( L_b * ctx_b + L_b * (L_b + 1) / 2 )   (Q.K^T and P.V over the unmasked pairs)
"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vllm_splitwise_b200.prefix_prefill import context_attention_fwd  # noqa: E402

PEAKS = {}
try:
    PEAKS = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))
except Exception:
    pass


def build(B, H, Hkv, D, bs, L, ctx, dtype, dev="cuda"):
    g = torch.Generator(device=dev).manual_seed(0)
    scale = D ** -0.5
    n = B * L
    q = torch.empty(n, H, D, dtype=dtype, device=dev).uniform_(-scale, scale, generator=g)
    k = torch.empty(n, Hkv, D, dtype=dtype, device=dev).uniform_(-scale, scale, generator=g)
    v = torch.empty(n, Hkv, D, dtype=dtype, device=dev).uniform_(-scale, scale, generator=g)
    nb = max(1, (ctx + bs - 1) // bs)
    NB = B * nb
    kc = torch.empty(NB, Hkv, D // 8, bs, 8, dtype=dtype, device=dev).uniform_(-scale, scale, generator=g)
    vc = torch.empty(NB, Hkv, D, bs, dtype=dtype, device=dev).uniform_(-scale, scale, generator=g)
    bt = torch.randperm(NB, device=dev, generator=g).view(B, nb).int()
    start = (torch.arange(B, device=dev) * L).int()
    seq = torch.full((B,), L + ctx, dtype=torch.int32, device=dev)
    cl = torch.full((B,), ctx, dtype=torch.int32, device=dev)
    return q, k, v, kc, vc, bt, start, seq, cl


def time_ms(fn, iters, warmup=3):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def run(name, B, H, Hkv, D, bs, L, ctx, dtype, iters=20):
    q, k, v, kc, vc, bt, start, seq, cl = build(B, H, Hkv, D, bs, L, ctx, dtype)
    o = torch.empty_like(q)
    fn = lambda: context_attention_fwd(q, k, v, o, kc, vc, bt, start, seq, cl, L)  # noqa: E731
    ms = time_ms(fn, iters)
    flops = 4.0 * D * H * B * (L * ctx + L * (L + 1) / 2)
    # library baseline: dense gather of the prefix (not timed) + SDPA with an explicit bottom-right causal mask
    G = H // Hkv
    nb = bt.shape[1]
    kd = kc[bt.long()].permute(0, 1, 4, 2, 3, 5).reshape(B, nb * bs, Hkv, D)[:, :ctx]   # [B, ctx, Hkv, D]
    vd = vc[bt.long()].permute(0, 1, 4, 2, 3).reshape(B, nb * bs, Hkv, D)[:, :ctx]
    kk = torch.cat([kd, k.view(B, L, Hkv, D)], 1).transpose(1, 2)                        # [B, Hkv, ctx+L, D]
    vv = torch.cat([vd, v.view(B, L, Hkv, D)], 1).transpose(1, 2)
    qq = q.view(B, L, H, D).transpose(1, 2)
    mask = (torch.arange(ctx + L, device="cuda")[None, :] <= (ctx + torch.arange(L, device="cuda"))[:, None])
    sd = lambda: torch.nn.functional.scaled_dot_product_attention(qq, kk, vv, attn_mask=mask, enable_gqa=G > 1)  # noqa: E731
    try:
        ms_sdpa = time_ms(sd, max(3, iters // 4))
        ref = sd().transpose(1, 2).reshape(B * L, H, D)
        err = (o.float() - ref.float()).abs().max().item()
    except Exception as e:  # noqa: BLE001
        ms_sdpa, err = None, str(e)[:80]
    peak = PEAKS.get("bf16_tflops")
    tf = flops / ms / 1e9
    print(json.dumps({"case": name, "workload": f"B={B} H={H} Hkv={Hkv} D={D} bs={bs} new={L} ctx={ctx} {dtype}",
                      "flops": flops, "ms": round(ms, 4), "TFLOPs": round(tf, 1), "bound": "tensor",
                      "frac_of_measured_bf16_peak": round(tf / peak, 4) if peak else None,
                      "torch_sdpa_masked_ms": round(ms_sdpa, 4) if ms_sdpa else None,
                      "speedup_vs_torch_sdpa": round(ms_sdpa / ms, 2) if ms_sdpa else None,
                      "max_abs_diff_vs_sdpa": err}), flush=True)


CASES = [
    ("llama7b_chunk2k_ctx2k_f16", 8, 32, 32, 128, 16, 2048, 2048, torch.float16),
    ("llama7b_prefill4k_noctx_f16", 4, 32, 32, 128, 16, 4096, 0, torch.float16),
    ("llama70b_tp8_chunk1k_ctx3k_bf16", 16, 8, 1, 128, 16, 1024, 3072, torch.bfloat16),
    ("mistral_gqa4_chunk512_ctx7680_bf16", 8, 32, 8, 128, 16, 512, 7680, torch.bfloat16),
    ("ref_test_shape_f16", 10, 64, 8, 128, 32, 512, 512, torch.float16),
    ("llama7b_d64_chunk2k_ctx2k_f16", 8, 32, 32, 64, 16, 2048, 2048, torch.float16),
]

if __name__ == "__main__":
    only = sys.argv[1] if len(sys.argv) > 1 else None      # run one case (for ncu)
    iters = int(sys.argv[2]) if len(sys.argv) > 2 else 20
    for c in CASES:
        if only is None or c[0] == only:
            run(*c, iters=iters)
