#!/usr/bin/env python
"""Cache-side rows of the hot path measured on one B200: reshape_and_cache at the
C4 / C5 shapes, copy_blocks, swap_blocks (D2D / D2H / H2D) and convert_fp8_e5m2 -- each as algorithmic
GB/s against MEASURED_PEAKS.json hbm_gbs (PCIe rows: reported only), next to the UNMODIFIED reference
CUDA kernel (oracle/_ref) on identical inputs with a bit-for-bit comparison of the results.

    python tools/cache_ops_bench.py [--iters 50] > gpurun_out/cache_ops.jsonl

One JSON line per case.  Timing: CUDA events over `--iters` calls after 3 warm-ups; consecutive calls
rotate over enough distinct buffers that the bytes touched between two uses of a buffer exceed L2 (126 MB),
except for the decode-sized (latency) cases where that is stated in the line.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import torch  # noqa: E402


def peak():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        return 6650.0


NCU_MODE = False  # --ncu: one warm-up, no graph replays, no reference kernels (keeps the ncu launch list short)


def timeit(fn, iters, nbuf):
    for i in range(1 if NCU_MODE else 3):
        fn(i % nbuf)
    torch.cuda.synchronize()
    beg, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    beg.record()
    for i in range(iters):
        fn(i % nbuf)
    end.record()
    torch.cuda.synchronize()
    return beg.elapsed_time(end) / iters


def timeit_graph(fn, nbuf, calls=24, replays=20):
    """GPU time per call with the CPU out of the way: `calls` calls captured into one CUDA graph (the decode
    path is graph-captured in the reference, model_runner.py:826-834), replayed `replays` times."""
    if NCU_MODE:
        return None
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for i in range(3):
            fn(i % nbuf)
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for i in range(calls):
            fn(i % nbuf)
    g.replay()
    torch.cuda.synchronize()
    beg, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    beg.record()
    for _ in range(replays):
        g.replay()
    end.record()
    torch.cuda.synchronize()
    return beg.elapsed_time(end) / (replays * calls)


def emit(name, desc, nbytes, ms, ref_ms, exact, note="", bound="hbm", graph_ms=None, ref_graph_ms=None):
    pk = peak()
    line = {"case": name, "workload": desc, "algorithmic_bytes": nbytes, "ms": round(ms, 5),
            "GBps": round(nbytes / ms / 1e6, 1), "bound": bound,
            "frac_of_measured_hbm": round(nbytes / ms / 1e6 / pk, 4) if bound == "hbm" else None,
            "reference_cuda_ms": round(ref_ms, 5) if ref_ms else None,
            "speedup_vs_reference_cuda": round(ref_ms / ms, 2) if ref_ms else None,
            "bit_exact_vs_reference_cuda": exact, "note": note}
    if graph_ms is not None:
        line["graph_replay_ms"] = round(graph_ms, 5)
        line["graph_replay_GBps"] = round(nbytes / graph_ms / 1e6, 1)
        if bound == "hbm":
            line["graph_replay_frac_of_measured_hbm"] = round(nbytes / graph_ms / 1e6 / pk, 4)
        if ref_graph_ms:
            line["reference_cuda_graph_replay_ms"] = round(ref_graph_ms, 5)
            line["graph_replay_speedup_vs_reference_cuda"] = round(ref_graph_ms / graph_ms, 2)
    print(json.dumps(line), flush=True)


def bench_reshape(name, desc, N, Hq, Hkv, D, bs, dtype, kv, iters, ours, ref, contiguous_slots):
    dev = "cuda"
    es = 2
    ce = 1 if kv == "fp8_e5m2" else es
    x = 16 // ce
    cdt = torch.uint8 if kv == "fp8_e5m2" else dtype
    NB = ((N + bs - 1) // bs if contiguous_slots else N) + 8
    per_call = N * Hkv * D * 2 * (es + ce) + N * 8
    nbuf = max(2, min(8, int(300e6 // max(per_call, 1)) + 1))
    g = torch.Generator(device="cpu").manual_seed(0)
    scale = D ** -0.5
    qkvs, kcs, vcs = [], [], []
    for _ in range(nbuf):
        qkvs.append(torch.empty(N, Hq + 2 * Hkv, D, dtype=dtype, device=dev).uniform_(-scale, scale))
        kcs.append(torch.zeros(NB, Hkv, D // x, bs, x, dtype=cdt, device=dev))
        vcs.append(torch.zeros(NB, Hkv, D, bs, dtype=cdt, device=dev))
    if contiguous_slots:   # prefill: a prompt's tokens fill consecutive slots of its blocks, blocks permuted
        blocks = torch.randperm(NB, generator=g)[:(N + bs - 1) // bs]
        slots = (blocks[:, None] * bs + torch.arange(bs)[None, :]).reshape(-1)[:N].contiguous()
    else:                  # decode: one token per sequence, each in a different block, random offset
        blocks = torch.randperm(NB, generator=g)[:N]
        slots = blocks * bs + torch.randint(0, bs, (N,), generator=g)
    slots = slots.to(torch.int64).to(dev)

    def call(mod, i):
        k = qkvs[i][:, Hq:Hq + Hkv]          # strided views of the fused qkv buffer, as llama.py:147-160 makes them
        v = qkvs[i][:, Hq + Hkv:]
        mod.reshape_and_cache(k, v, kcs[i], vcs[i], slots, kv)

    views = [(q[:, Hq:Hq + Hkv], q[:, Hq + Hkv:]) for q in qkvs]   # made once: view construction is not the op

    def call(mod, i):  # noqa: F811
        mod.reshape_and_cache(views[i][0], views[i][1], kcs[i], vcs[i], slots, kv)

    ms = timeit(lambda i: call(ours, i), iters, nbuf)
    gms = timeit_graph(lambda i: call(ours, i), nbuf)
    mine = (kcs[0].clone(), vcs[0].clone())
    ref_ms, exact, rgms = None, None, None
    if ref is not None:
        kcs[0].zero_(); vcs[0].zero_()
        ref_ms = timeit(lambda i: call(ref.cache_ops, i), iters, nbuf)
        rgms = timeit_graph(lambda i: call(ref.cache_ops, i), nbuf)
        exact = bool(torch.equal(mine[0].view(torch.uint8), kcs[0].view(torch.uint8)) and
                     torch.equal(mine[1].view(torch.uint8), vcs[0].view(torch.uint8)))
    emit(name, desc, per_call, ms, ref_ms, exact, graph_ms=gms, ref_graph_ms=rgms,
         note=f"{nbuf} buffer sets rotated; K/V are strided views of a [N, Hq+2Hkv, D] qkv buffer; ms = eager calls "
              f"(CPU launch rate included), graph_replay_ms = the same calls replayed from a CUDA graph")


def bench_copy_blocks(iters, ours, ref):
    # C2-like: 32 layers, Hkv=32, D=128, bs=16 fp16 -> 128 KiB per block per K/V; 64 copy-on-write pairs
    L, Hkv, D, bs, NB, pairs = 32, 32, 128, 16, 512, 64
    dev = "cuda"
    kc = [torch.empty(NB, Hkv, D // 8, bs, 8, dtype=torch.half, device=dev).uniform_(-1, 1) for _ in range(L)]
    vc = [torch.empty(NB, Hkv, D, bs, dtype=torch.half, device=dev).uniform_(-1, 1) for _ in range(L)]
    g = torch.Generator().manual_seed(0)
    maps = []
    for r in range(4):   # 4 disjoint (src, dst) sets rotated: 4 x 1.07 GB touched between reuses
        perm = torch.randperm(NB, generator=g)[:2 * pairs].tolist()
        maps.append({perm[2 * i]: [perm[2 * i + 1]] for i in range(pairs)})
    nbytes = 2 * L * pairs * 2 * (Hkv * D * bs * 2)
    ms = timeit(lambda i: ours.copy_blocks(kc, vc, maps[i]), iters, 4)
    snap = [t.clone() for t in kc[:2] + vc[:2]]
    ref_ms, exact = None, None
    if ref is not None:
        ref_ms = timeit(lambda i: ref.cache_ops.copy_blocks(kc, vc, maps[i]), iters, 4)
        exact = all(torch.equal(a, b) for a, b in zip(snap, kc[:2] + vc[:2]))   # idempotent: same result either way
    emit("copy_blocks", f"{L} layers x {pairs} pairs x (K,V) x 128 KiB blocks, fp16 (C2 cache shape)", nbytes, ms,
         ref_ms, exact, note="bytes = read + write; reference time includes its synchronous table upload")


def bench_swap(iters, ours, ref):
    Hkv, D, bs, NB, n = 32, 128, 16, 1024, 256
    dev = "cuda"
    g = torch.Generator().manual_seed(0)
    src = torch.empty(NB, Hkv, D, bs, dtype=torch.half, device=dev).uniform_(-1, 1)
    dst = torch.zeros_like(src)
    host = torch.zeros(NB, Hkv, D, bs, dtype=torch.half).pin_memory()
    bb = Hkv * D * bs * 2
    for pattern in ("fragmented", "contiguous"):
        if pattern == "fragmented":
            s, d = torch.randperm(NB, generator=g)[:n].tolist(), torch.randperm(NB, generator=g)[:n].tolist()
        else:
            s, d = list(range(100, 100 + n)), list(range(300, 300 + n))
        mapping = dict(zip(s, d))
        for kind, a, b, bound in (("d2d", src, dst, "hbm"), ("d2h", src, host, "pcie"), ("h2d", host, dst, "pcie")):
            nbytes = n * bb * (2 if kind == "d2d" else 1)
            it = iters if kind == "d2d" else max(5, iters // 5)
            b.zero_()
            ms = timeit(lambda i: ours.swap_blocks(a, b, mapping), it, 1)
            # D2D: the eager call is bound by the host side of the op (a 256-entry dict crosses the binding per call);
            # the replayed graph shows the gather kernel itself
            gms = timeit_graph(lambda i: ours.swap_blocks(a, b, mapping), 1) if kind == "d2d" else None
            mine = b.clone() if kind != "d2h" else b.clone()
            ref_ms, exact = None, None
            if ref is not None:
                b.zero_()
                ref_ms = timeit(lambda i: ref.cache_ops.swap_blocks(a, b, mapping), it, 1)
                exact = bool(torch.equal(mine, b))
            emit(f"swap_blocks_{kind}_{pattern}", f"{n} blocks x 128 KiB, one layer's V, fp16", nbytes, ms, ref_ms, exact,
                 note="bytes = read + write; ms = eager op calls (host-bound), graph_replay_ms = the gather kernel"
                 if kind == "d2d" else "bytes = one direction over PCIe (pinned host)", bound=bound, graph_ms=gms)


def bench_convert(iters, ours, ref):
    n = 1 << 28   # 256 Mi elements: 512 MB fp16 -> 256 MB fp8
    src = torch.empty(n, dtype=torch.half, device="cuda").uniform_(-1, 1)
    dst = torch.empty(n, dtype=torch.uint8, device="cuda")
    back = torch.empty_like(src)
    ms = timeit(lambda i: ours.convert_fp8_e5m2(src, dst), iters, 1)
    mine = dst.clone()
    ref_ms, exact = None, None
    if ref is not None:
        ref_ms = timeit(lambda i: ref.cache_ops.convert_fp8_e5m2(src, dst), iters, 1)
        exact = bool(torch.equal(mine, dst))
    emit("convert_fp16_to_fp8", "256 Mi elements", n * 3, ms, ref_ms, exact)
    ms = timeit(lambda i: ours.convert_fp8_e5m2(dst, back), iters, 1)
    mine = back.clone()
    if ref is not None:
        ref_ms = timeit(lambda i: ref.cache_ops.convert_fp8_e5m2(dst, back), iters, 1)
        exact = bool(torch.equal(mine.view(torch.int16), back.view(torch.int16)))
    emit("convert_fp8_to_fp16", "256 Mi elements", n * 3, ms, ref_ms, exact)


def bench_rope(name, desc, N, Hq, Hkv, D, bs, dtype, kv, ours_ops, ours_cache, ref):
    """The decode step's rotary_embedding + reshape_and_cache pair (llama.py:156-158 -> attention backend):
    ours fused (one launch), ours as two ops, the reference's two kernels; CUDA-graph replay times."""
    sys.path.insert(0, ROOT)
    from oracle import paged_kv_oracle as O   # cos/sin table recipe only (a checker-side helper, not timed)
    dev = "cuda"
    ce = 1 if kv == "fp8_e5m2" else 2
    x = 16 // ce
    cdt = torch.uint8 if kv == "fp8_e5m2" else dtype
    NB = N + 8
    nbuf = 4
    g = torch.Generator().manual_seed(0)
    cache = O.compute_cos_sin_cache(D, 8192, 10000, dtype).to(dev)
    pos = torch.randint(0, 8192, (N,), generator=g).to(dev)
    slots = (torch.randperm(NB, generator=g)[:N] * bs + torch.randint(0, bs, (N,), generator=g)).to(dev)
    qkvs = [torch.randn(N, (Hq + 2 * Hkv) * D, dtype=dtype, device=dev) for _ in range(nbuf)]
    views = [(q[:, :Hq * D], q[:, Hq * D:(Hq + Hkv) * D], q[:, (Hq + Hkv) * D:]) for q in qkvs]
    kvs3 = [(k.view(N, Hkv, D), v.view(N, Hkv, D)) for _, k, v in views]
    kcs = [torch.zeros(NB, Hkv, D // x, bs, x, dtype=cdt, device=dev) for _ in range(nbuf)]
    vcs = [torch.zeros(NB, Hkv, D, bs, dtype=cdt, device=dev) for _ in range(nbuf)]

    def fused(i):
        ours_cache.rotary_embedding_reshape_and_cache(pos, views[i][0], views[i][1], views[i][2], D, cache, True,
                                                      kcs[i], vcs[i], slots, kv)

    def two(ops_mod, cache_mod):
        def f(i):
            ops_mod.rotary_embedding(pos, views[i][0], views[i][1], D, cache, True)
            cache_mod.reshape_and_cache(kvs3[i][0], kvs3[i][1], kcs[i], vcs[i], slots, kv)
        return f

    t_f = timeit_graph(fused, nbuf)
    t_2 = timeit_graph(two(ours_ops, ours_cache), nbuf)
    t_r = timeit_graph(two(ref.ops, ref.cache_ops), nbuf) if ref is not None else None
    es = 2
    nbytes = N * (2 * (Hq + Hkv) * D * es + Hkv * D * es + 2 * Hkv * D * ce) + N * 16
    line = {"case": name, "workload": desc, "algorithmic_bytes": nbytes, "fused_graph_replay_ms": round(t_f, 5),
            "two_ops_graph_replay_ms": round(t_2, 5),
            "reference_cuda_two_kernels_graph_replay_ms": round(t_r, 5) if t_r else None,
            "fused_speedup_vs_reference_cuda": round(t_r / t_f, 2) if t_r else None,
            "note": "latency case: per-call GPU time of 24 calls captured in one CUDA graph; bit-exactness of the "
                    "fused op against both sequences is asserted in tests/test_rope_gpu.py"}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=50)
    ap.add_argument("--only", default="")
    ap.add_argument("--ncu", action="store_true", help="short launch list for an ncu capture: our kernels only")
    args = ap.parse_args()
    assert torch.cuda.is_available()
    from vllm_splitwise_b200._C import cache_ops as ours
    from _util import load_ref_cuda
    global NCU_MODE
    NCU_MODE = args.ncu
    ref = None if args.ncu else load_ref_cuda()
    only = set(filter(None, args.only.split(",")))

    def want(n):
        return not only or n in only

    if want("reshape"):
        bench_reshape("reshape_and_cache_c4_decode", "C4 (i): N=128 tokens, 8 kv heads, head 128, bf16 -> fp8_e5m2 (latency case)",
                      128, 32, 8, 128, 16, torch.bfloat16, "fp8_e5m2", args.iters * 4, ours, ref, False)
        bench_reshape("reshape_and_cache_c4_prefill", "C4 (ii): N=32768 tokens, 8 kv heads, head 128, bf16 -> fp8_e5m2",
                      32768, 32, 8, 128, 16, torch.bfloat16, "fp8_e5m2", args.iters, ours, ref, True)
        bench_reshape("reshape_and_cache_c4_prefill_bf16", "C4 (ii) with a bf16 cache: N=32768 tokens, 8 kv heads, head 128",
                      32768, 32, 8, 128, 16, torch.bfloat16, "auto", args.iters, ours, ref, True)
        bench_reshape("reshape_and_cache_c5_prefill", "C5 prefill: N=2048 tokens, 10 kv heads (tp=4 shard), head 128, fp16",
                      2048, 10, 10, 128, 16, torch.float16, "auto", args.iters * 2, ours, ref, True)
        bench_reshape("reshape_and_cache_c2_decode", "C2 decode: N=64 tokens, 32 kv heads, head 128, fp16 (latency case)",
                      64, 32, 32, 128, 16, torch.float16, "auto", args.iters * 4, ours, ref, False)
        bench_reshape("reshape_and_cache_c3_decode", "C3 tp=8 decode: N=256 tokens, 1 kv head, head 128, bf16 (latency case)",
                      256, 8, 1, 128, 16, torch.bfloat16, "auto", args.iters * 4, ours, ref, False)
    if want("rope"):
        from vllm_splitwise_b200._C import ops as ours_ops
        bench_rope("rope_reshape_and_cache_c4_decode", "C4 decode step: N=128 tokens, 32 q / 8 kv heads, head 128, bf16 -> fp8_e5m2",
                   128, 32, 8, 128, 16, torch.bfloat16, "fp8_e5m2", ours_ops, ours, ref)
        bench_rope("rope_reshape_and_cache_c2_decode", "C2 decode step: N=64 tokens, 32 q / 32 kv heads, head 128, fp16",
                   64, 32, 32, 128, 16, torch.float16, "auto", ours_ops, ours, ref)
        bench_rope("rope_reshape_and_cache_c3_decode", "C3 tp=8 decode step: N=256 tokens, 8 q / 1 kv head, head 128, bf16",
                   256, 8, 1, 128, 16, torch.bfloat16, "auto", ours_ops, ours, ref)
    if want("copy"):
        bench_copy_blocks(args.iters, ours, ref)
    if want("swap"):
        bench_swap(args.iters, ours, ref)
    if want("convert"):
        bench_convert(max(5, args.iters // 5), ours, ref)


if __name__ == "__main__":
    main()
