"""Full BASELINE.json sizes, ours next to the UNMODIFIED reference CUDA kernels (oracle/_ref) on identical
inputs: (1) parity of the outputs at full size, (2) both kernels' device time with CUDA events, written to
gpurun_out/ref_cuda_compare.json (a record, not an assertion -- the only perf assertion is "ours is not
slower").  Skipped when oracle/_ref was not built.

The reference benchmark's own recipe (benchmarks/kernels/benchmark_paged_attention.py:143-153) times
back-to-back launches; here inputs are distinct physical blocks of a cache larger than L2, two caches
alternated, CUDA events instead of perf_counter.
"""
import json
import os

import pytest
import torch

from _util import ROOT, load_ref_cuda, make_attention_inputs, run_v1, run_v2

pytestmark = pytest.mark.gpu

CASES = {
    "C2_llama7b_f16": dict(S=64, heads=(32, 32), ctx=4096, dtype=torch.half, kv="auto"),
    "C3_llama70b_tp8_bf16": dict(S=256, heads=(8, 1), ctx=4096, dtype=torch.bfloat16, kv="auto"),
    "C3_llama70b_tp2_bf16": dict(S=256, heads=(32, 4), ctx=4096, dtype=torch.bfloat16, kv="auto"),
    "C4_mistral_fp8": dict(S=128, heads=(32, 8), ctx=8192, dtype=torch.bfloat16, kv="fp8_e5m2"),
}
RESULTS = {}


def _time(fn, n=20):
    for _ in range(3):
        fn(0)
        fn(1)
    torch.cuda.synchronize()
    beg, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    beg.record()
    for i in range(n):
        fn(i & 1)
    end.record()
    torch.cuda.synchronize()
    return beg.elapsed_time(end) / n


@pytest.mark.parametrize("name", list(CASES))
def test_full_size_vs_reference_kernel(name):
    ref = load_ref_cuda()
    if ref is None:
        pytest.skip("oracle/_ref/vllm_ref_C.so not built")
    from vllm_splitwise_b200._C import ops
    c = CASES[name]
    S, ctx = c["S"], c["ctx"]
    nb = S * (ctx // 16)
    pair = [make_attention_inputs(num_seqs=S, num_heads=c["heads"], head_size=128, block_size=16, dtype=c["dtype"],
                                  kv_cache_dtype=c["kv"], num_blocks=nb, context_lens=[ctx] * S, distinct_blocks=True,
                                  seed=s) for s in (0, 1)]
    atol = 1e-2 if c["kv"] == "fp8_e5m2" else 1e-3
    rec = {}
    for version, run in (("v1", lambda o, a: run_v1(o, a)), ("v2", lambda o, a: run_v2(o, a)[0])):
        mine, theirs = run(ops, pair[0]), run(ref.ops, pair[0])
        torch.cuda.synchronize()
        err = (mine.float() - theirs.float()).abs().max().item()
        assert torch.allclose(mine.float(), theirs.float(), atol=atol, rtol=1e-3), f"{name} {version}: max err {err}"
        t_mine = _time(lambda i: run(ops, pair[i]))
        t_ref = _time(lambda i: run(ref.ops, pair[i]))
        rec[version] = {"ours_ms": round(t_mine, 4), "reference_cuda_ms": round(t_ref, 4),
                        "speedup": round(t_ref / t_mine, 3), "max_abs_diff": err}
        assert t_mine <= t_ref * 1.05, f"{name} {version}: ours {t_mine:.3f} ms vs reference kernel {t_ref:.3f} ms"
    RESULTS[name] = rec
    out_dir = os.path.join(ROOT, "gpurun_out")
    os.makedirs(out_dir, exist_ok=True)
    with open(os.path.join(out_dir, "ref_cuda_compare.json"), "w") as f:
        json.dump(RESULTS, f, indent=1)
