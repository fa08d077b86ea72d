#!/usr/bin/env python
"""Bisection helper: run one test_paged_attention case repeatedly, print where the error sits."""
import os, sys, random
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
from _util import make_attention_inputs, oracle_attention, run_v1, run_v2
from vllm_splitwise_b200._C import ops

def case(version, num_heads, head_size, use_alibi, block_size, dtype, kv, reps=4):
    random.seed(0)
    MAX = 4096 + 333
    ctx = [random.randint(1, MAX) for _ in range(7)]; ctx[-1] = MAX
    a = make_attention_inputs(num_seqs=7, num_heads=num_heads, head_size=head_size, block_size=block_size, dtype=dtype,
                              kv_cache_dtype=kv, num_blocks=4321, context_lens=ctx, use_alibi=use_alibi)
    ref = oracle_attention(a).float()
    for r in range(reps):
        out = (run_v1(ops, a) if version == "v1" else run_v2(ops, a)[0]).float()
        torch.cuda.synchronize()
        err = (out - ref).abs()
        per = err.amax(dim=2)  # [S, H]
        bad = (per > 2e-3).nonzero().tolist()
        print(f"{version} {num_heads} D={head_size} alibi={use_alibi} bs={block_size} {dtype} {kv} rep{r}: max {err.max().item():.4g} "
              f"nan={int(torch.isnan(out).sum())} bad(seq,head)={bad[:12]} ctx={ctx}", flush=True)
        if bad and r == 0 and a["alibi"] is not None:
            print("  alibi slopes of bad heads:", [round(a["alibi"][h].item(), 3) for _, h in bad[:12]])

if __name__ == "__main__":
    bf, hf = torch.bfloat16, torch.half
    case("v2", (64, 8), 128, True, 16, bf, "auto")
    case("v1", (64, 8), 128, True, 16, bf, "auto")
    case("v2", (64, 8), 128, True, 16, hf, "auto", 2)
    case("v2", (64, 8), 128, True, 32, bf, "auto", 2)
    case("v2", (64, 8), 256, True, 16, bf, "auto", 2)
    os.environ["B200_ATTN_FORCE_SIMT"] = "1"
