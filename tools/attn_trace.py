"""Development tool: timeline of CTA 0 of the decode kernel in quad mode (v2 signature, tensor-core consumer).

Needs a library built with `make -C vllm_splitwise_b200/csrc TRACE=1`.  Per quad (the CTA's k-th work unit) every
consumer warp stamps: 0 quad known, 1 query fragments ready, 2 first block landed, 5 second block landed,
3 last block done, 4 partial stored.      python tools/attn_trace.py
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

trace = torch.zeros(4 * 32 * 8, dtype=torch.int64, device="cuda")
os.environ["B200_ATTN_TRACE_PTR"] = hex(trace.data_ptr())

from _util import make_attention_inputs  # noqa: E402
from vllm_splitwise_b200._C import ops  # noqa: E402


def main():
    S, ctx = 128, 8192   # C4: Mistral-7B GQA, bf16 query, fp8 KV
    a = make_attention_inputs(num_seqs=S, num_heads=(32, 8), head_size=128, block_size=16, dtype=torch.bfloat16,
                              kv_cache_dtype="fp8_e5m2", num_blocks=S * ctx // 16, context_lens=[ctx] * S, seed=0,
                              distinct_blocks=True)
    q = a["q"]
    P = (ctx + 511) // 512
    out = torch.empty_like(q)
    tmp = torch.empty(S, 32, P, 128, dtype=q.dtype, device="cuda")
    es = torch.empty(S, 32, P, dtype=torch.float32, device="cuda")
    ml = torch.empty_like(es)
    for _ in range(3):
        trace.zero_()
        ops.paged_attention_v2(out, es, ml, tmp, q, a["kc"], a["vc"], a["hkv"], a["scale"], a["bt"], a["cl"], 16, ctx,
                               None, "fp8_e5m2")
        torch.cuda.synchronize()
    t = trace.cpu().view(4, 32, 8)
    if int(t.max()) == 0:
        print("no stamps: not a TRACE=1 build, or the call did not take the quad mode")
        return
    t0 = int(t[t > 0].min())
    n = max(k for k in range(32) if int(t[:, k].max()) > 0) + 1
    print("quad | warp: known qfrag first second last stored")
    for k in range(min(n, 8)):
        for w in range(4):
            print(f"{k:3d} w{w}: " + " ".join(f"{int(t[w, k, e]) - t0:8d}" if int(t[w, k, e]) else "       -" for e in (0, 1, 2, 5, 3, 4)))

    def avg(e1, e2, shift=0):
        d = [int(t[w, k + shift, e2]) - int(t[w, k, e1]) for w in range(4) for k in range(1, n - 1 - shift)
             if int(t[w, k, e1]) and int(t[w, k + shift, e2])]
        return sum(d) / max(1, len(d))
    print(f"quads seen {n}; per quad: period {avg(0, 0, 1):.0f}  decode+query fragments {avg(0, 1):.0f}  wait first block {avg(1, 2):.0f}  "
          f"first->second block {avg(2, 5):.0f}  blocks (first landed -> last done) {avg(2, 3):.0f}  store partial {avg(3, 4):.0f}  "
          f"stored -> next quad known {avg(4, 0, 1):.0f}")


if __name__ == "__main__":
    main()
