"""Development tool: phase timeline of ONE CTA of the 256-row prefix-prefill kernel.

Needs a library built with `make -C vllm_splitwise_b200/csrc TRACE=1` (clock64 stamps compiled in; the product build has
none).  CTA (0,0,0) -- the heaviest row block of the first (head, sequence) pair -- stamps, per KV tile:
  softmax warps 0-7 (0-3 block A, 4-7 block B): 0 S ready, 1 S in registers (tcgen05.ld done), 2 maximum known,
                                                3 P stored, 4 P announced
  MMA warp (8):   0 P_A seen, 1 P_A.V + next Q_A.K^T issued, 2 P_B seen, 3 P_B.V + next Q_B.K^T issued,
                  4 V tile seen (block A's wait), 6 V tile seen (block B's wait), 5 / 7 P.V issued (A / B), about to wait for K
  V loader (9):   0 stage free, 1 copies issued, 2 tile announced         K loader (10): the same
Prints cycles relative to the first stamp for a few steady-state tiles and the per-phase averages.
    python tools/ctx_trace.py [case-name]
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

NWHO = 11
NT = 64
trace = torch.zeros(NWHO * NT * 8, dtype=torch.int64, device="cuda")
os.environ["B200_CTX_TRACE_PTR"] = hex(trace.data_ptr())

from prefix_prefill_bench import CASES, build  # noqa: E402
from vllm_splitwise_b200.prefix_prefill import context_attention_fwd  # noqa: E402


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "llama7b_chunk2k_ctx2k_f16"
    case = [c for c in CASES if c[0] == name][0]
    _, B, H, Hkv, D, bs, L, ctx, dtype = case
    q, k, v, kc, vc, bt, start, seq, cl = build(B, H, Hkv, D, bs, L, ctx, dtype)
    o = torch.empty_like(q)
    for _ in range(3):
        trace.zero_()
        context_attention_fwd(q, k, v, o, kc, vc, bt, start, seq, cl, L)
        torch.cuda.synchronize()
    t = trace.cpu().view(NWHO, NT, 8)
    if int(t.max()) == 0:
        print("no stamps: the library was not built with TRACE=1")
        return
    t0 = int(t[t > 0].min())
    n = max(i for i in range(NT) if int(t[:, i].max()) > 0) + 1

    def rel(w, i, e):
        return int(t[w, i, e]) - t0 if int(t[w, i, e]) else None

    def fmt(x):
        return f"{x:7d}" if x is not None else "      -"
    for i in [j for j in (n // 4, n // 4 + 1, n // 4 + 2, 3 * n // 4 - 2, 3 * n // 4 - 1, 3 * n // 4) if 0 <= j < n - 1]:
        print(f"--- tile {i}")
        for w in range(8):
            print(f"  softmax warp {w} ({'AB'[w >> 2]}): " + " ".join(fmt(rel(w, i, e)) for e in range(5)))
        print("  mma: V(A) " + fmt(rel(8, i, 4)) + " P_A " + fmt(rel(8, i, 0)) + " PV_A issued " + fmt(rel(8, i, 5)) + " issuedA " + fmt(rel(8, i, 1)) +
              " V(B) " + fmt(rel(8, i, 6)) + " P_B " + fmt(rel(8, i, 2)) + " PV_B issued " + fmt(rel(8, i, 7)) + " issuedB " + fmt(rel(8, i, 3)))
        print("  V loader: free " + fmt(rel(9, i, 0)) + " issued " + fmt(rel(9, i, 1)) + " announced " + fmt(rel(9, i, 2)))
        print("  K loader: free " + fmt(rel(10, i, 0)) + " issued " + fmt(rel(10, i, 1)) + " announced " + fmt(rel(10, i, 2)))
    lo, hi = 4, max(5, n - 4)   # steady state: skip the first and last tiles

    def avg(w1, e1, w2, e2, shift=0):
        d = [int(t[w2, j + shift, e2]) - int(t[w1, j, e1]) for j in range(lo, hi - shift)
             if int(t[w1, j, e1]) and int(t[w2, j + shift, e2])]
        return sum(d) / max(1, len(d))
    print("per softmax warp: ld / max / exp+st / announce / (announce -> next S ready)")
    for w in range(8):
        print(f"  warp {w}: {avg(w, 0, w, 1):6.0f} {avg(w, 1, w, 2):6.0f} {avg(w, 2, w, 3):6.0f} {avg(w, 3, w, 4):6.0f} {avg(w, 4, w, 0, 1):6.0f}"
              f"   period {avg(w, 0, w, 0, 1):6.0f}")
    lastA = [max(int(t[w, j, 4]) for w in range(4)) for j in range(NT)]
    lastB = [max(int(t[w, j, 4]) for w in range(4, 8)) for j in range(NT)]
    dA = [int(t[8, j, 0]) - lastA[j] for j in range(lo, hi) if lastA[j] and int(t[8, j, 0])]
    dB = [int(t[8, j, 2]) - lastB[j] for j in range(lo, hi) if lastB[j] and int(t[8, j, 2])]
    print(f"last warp of A announces -> MMA sees P_A: {sum(dA) / max(1, len(dA)):.0f}   B: {sum(dB) / max(1, len(dB)):.0f}")
    print(f"mma: period {avg(8, 0, 8, 0, 1):.0f}  [P_A seen -> A issued] {avg(8, 0, 8, 1):.0f}  [-> V(B) seen] {avg(8, 1, 8, 6):.0f}  "
          f"[-> P_B seen] {avg(8, 6, 8, 2):.0f}  [-> B issued] {avg(8, 2, 8, 3):.0f}  [-> next V(A) seen] {avg(8, 3, 8, 4, 1):.0f}  "
          f"[-> next P_A seen] {avg(8, 4, 8, 0, 1):.0f}")
    print(f"A issued -> S_A ready (warp 0) {avg(8, 1, 0, 0, 1):.0f}   B issued -> S_B ready (warp 4) {avg(8, 3, 4, 0, 1):.0f}")
    print(f"V loader: free -> issued {avg(9, 0, 9, 1):.0f}  issued -> announced {avg(9, 1, 9, 2):.0f}  announced -> MMA needs it (A) {avg(9, 2, 8, 4):.0f}")
    print(f"K loader: free -> issued {avg(10, 0, 10, 1):.0f}  issued -> announced {avg(10, 1, 10, 2):.0f}")


if __name__ == "__main__":
    main()
