#!/usr/bin/env python
"""Run paged_attention a few times on one workload shape (short command for ncu / compute-sanitizer / bisection).

    python tools/attn_once.py --workload c2 --version v1 --iters 3 [--S 64 --ctx 4096] [--check]
"""
import argparse
import math
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import torch  # noqa: E402

import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="c2")
    ap.add_argument("--version", default="v1")
    ap.add_argument("--iters", type=int, default=3)
    ap.add_argument("--S", type=int)
    ap.add_argument("--ctx", type=int)
    ap.add_argument("--check", action="store_true", help="compare 3 sequences with the oracle")
    ap.add_argument("--rc", action="store_true", help="also run reshape_and_cache each iteration")
    args = ap.parse_args()
    w = dict(bench.WORKLOADS[args.workload])
    if args.S:
        w["S"] = args.S
    if args.ctx:
        w["ctx"] = args.ctx
    from vllm_splitwise_b200._C import cache_ops, ops
    dev = torch.device("cuda:0")
    wl = bench.build_device_workload(w, 1, dev, 0)
    d = wl["dev"]
    out = torch.empty_like(d["q"])
    P = math.ceil(w["ctx"] / 512)
    tmp = torch.empty(w["S"], w["Hq"], P, w["D"], dtype=out.dtype, device=dev)
    es = torch.empty(w["S"], w["Hq"], P, dtype=torch.float32, device=dev)
    ml = torch.empty_like(es)
    torch.cuda.synchronize()
    for _ in range(args.iters):
        if args.rc:
            cache_ops.reshape_and_cache(d["key"], d["value"], wl["kcs"][0], wl["vcs"][0], d["slots"], w["kv"])
        if args.version == "v1":
            ops.paged_attention_v1(out, d["q"], wl["kcs"][0], wl["vcs"][0], w["Hkv"], wl["scale"], d["bt"], d["cl"],
                                   w["bs"], w["ctx"], None, w["kv"])
        else:
            ops.paged_attention_v2(out, es, ml, tmp, d["q"], wl["kcs"][0], wl["vcs"][0], w["Hkv"], wl["scale"],
                                   d["bt"], d["cl"], w["bs"], w["ctx"], None, w["kv"])
    torch.cuda.synchronize()
    msg = f"ok {args.workload} {args.version} S={w['S']} ctx={w['ctx']}"
    if args.check:
        from _util import oracle_attention
        idx = [0, w["S"] // 2, w["S"] - 1]
        a = dict(q=d["q"][idx], kc=wl["kcs"][0], vc=wl["vcs"][0], bt=d["bt"][idx], cl=d["cl"][idx], alibi=None,
                 scale=wl["scale"], hq=w["Hq"], hkv=w["Hkv"], kv_cache_dtype=w["kv"])
        ref = oracle_attention(a)
        msg += f" max_err={(out[idx].float() - ref.float()).abs().max().item():.3e}"
    print(msg, flush=True)


if __name__ == "__main__":
    main()
