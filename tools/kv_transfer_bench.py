#!/usr/bin/env python
"""C5: Splitwise prefill -> decode KV-block transfer over NVLink, one process per GPU.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        tools/kv_transfer_bench.py [--prompts 8] [--layers 40] [--iters 5]

N even: ranks [0, N/2) are prompt GPUs, rank r sends to token rank r + N/2 (reference comm_utils.py:165;
N=8 is BASELINE config 5: 4 prefill -> 4 decode GPUs).  Shape = Llama-2-13B tp=4 shard: 40 layers, 10 KV
heads/GPU, head 128, block 16, fp16 -> block_bytes 40 960; a 2048-token prompt = 128 blocks =
419 430 400 B per GPU pair.

For every (block-list pattern, mode) it times `--iters` rounds of `--prompts` prompts in flight and
verifies on the token GPU, bit for bit, that exactly the transferred blocks arrived (checksum of int64
words against the value the prompt side computed on its source blocks, untouched blocks still zero).
Timing: CUDA events per rank from a common barrier to (prompt) the end of the signal kernel and (token)
the end of the wait kernel; the MAX over the ranks of a pair set is the transfer time.
Prints one JSON line per (pattern, mode) on rank 0.

modes   bulk       one put_layers launch per prompt (all layers, all runs) + signal
        per_layer  one put_runs launch per layer per prompt (how the attention backends call it while
                   prefill is still computing, flash_attn.py:81-86) + signal
        ref_api    exactly the reference call pattern: for every layer, for every prompt (semaphore), for every
                   coalesced run: put(sem, layer, block_start, num_blocks); after the last layer one
                   signal_and_flush(sem) per prompt (worker.py:327-331).  The manager batches the deferred
                   puts of a layer into one launch.
        paced      per_layer on the side stream while the main stream runs a stand-in for each layer's prefill compute
                   (a spin kernel of --pace-us per layer): reports time_to_signal_ms = end of the last layer's
                   compute -> signal delivered, i.e. what the decode side waits for after prefill has finished
        nccl       baseline: torch.distributed batch_isend_irecv, ONE batch per prompt over all layers and runs
                   (one P2P op per (layer, K|V, run)), no host wait before the end of the round

--sweep "variant:ctas:layers_per_launch,..." repeats every (pattern, mode) for each put configuration
(variant 0 = SIMT warp chunks, 1 = cp.async.bulk ring; csrc/kv_transfer.cu).
"""
import argparse
import json
import os
import random
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

NVLINK_MEASURED_GBS = 770.0   # peer copy per direction, /opt/skills/guides/B200_PROFILING.md
NVLINK_NOMINAL_GBS = 900.0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--prompts", type=int, default=8, help="prompts in flight per round")
    ap.add_argument("--layers", type=int, default=40)
    ap.add_argument("--kv-heads", type=int, default=10)
    ap.add_argument("--head-size", type=int, default=128)
    ap.add_argument("--block-size", type=int, default=16)
    ap.add_argument("--prompt-tokens", type=int, default=2048)
    ap.add_argument("--iters", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--max-ctas", type=int, default=0)
    ap.add_argument("--modes", default="bulk,per_layer,ref_api,nccl")
    ap.add_argument("--patterns", default="contiguous,fragmented")
    ap.add_argument("--nccl-max-prompts", type=int, default=2, help="cap for the fragmented NCCL baseline")
    ap.add_argument("--sweep", default="", help="variant:ctas:layers_per_launch[,...]; empty = the library defaults")
    ap.add_argument("--comm-stream", type=int, default=1, help="1 = puts and signals ride a side stream")
    ap.add_argument("--pace-us", type=float, default=500.0, help="paced mode: stand-in compute per layer")
    args = ap.parse_args()

    import torch
    import torch.distributed as dist
    rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    boot = dist.new_group(backend="gloo")
    results = run(rank, world, dev, boot, prompts=args.prompts, layers=args.layers, kv_heads=args.kv_heads,
                  head_size=args.head_size, block_size=args.block_size, prompt_tokens=args.prompt_tokens,
                  iters=args.iters, warmup=args.warmup, max_ctas=args.max_ctas, modes=args.modes.split(","),
                  patterns=args.patterns.split(","), nccl_max_prompts=args.nccl_max_prompts, echo=True,
                  sweep=[tuple(int(x) for x in c.split(":")) for c in args.sweep.split(",") if c],
                  comm_stream=bool(args.comm_stream), pace_us=args.pace_us)
    dist.barrier()
    dist.destroy_process_group()
    return 0 if all(r["bit_exact"] for r in results) else 1


def run(rank, world, dev, boot, prompts=8, layers=40, kv_heads=10, head_size=128, block_size=16, prompt_tokens=2048,
        iters=5, warmup=2, max_ctas=0, modes=("bulk", "per_layer", "ref_api", "nccl"),
        patterns=("contiguous", "fragmented"), nccl_max_prompts=2, echo=False, sweep=(), comm_stream=True,
        pace_us=500.0):
    """The C5 measurement as a function (bench.py calls it for `extra.kv_transfer`): torch.distributed must be
    initialised with the NCCL backend, `boot` is a gloo group for the one-time handle exchange.  Returns the
    result lines (every rank gets the list; the numbers are the max over ranks)."""
    import argparse as _argparse
    import torch
    import torch.distributed as dist
    from vllm_splitwise_b200.comm_utils import KVCacheCommManager
    from vllm_splitwise_b200.utils import coalesce_blocks, kv_cache_shapes
    args = _argparse.Namespace(prompts=prompts, layers=layers, kv_heads=kv_heads, head_size=head_size,
                               block_size=block_size, prompt_tokens=prompt_tokens, iters=iters, warmup=warmup,
                               max_ctas=max_ctas, modes=",".join(modes), patterns=",".join(patterns),
                               nccl_max_prompts=nccl_max_prompts)
    assert world % 2 == 0, "needs an even number of ranks"
    npw = world // 2
    is_prompt = rank < npw
    peer = (rank + npw) % world

    L, H, D, bs = args.layers, args.kv_heads, args.head_size, args.block_size
    blocks_per_prompt = (args.prompt_tokens + bs - 1) // bs
    NB = args.prompts * blocks_per_prompt * 2  # room for non-adjacent fragmented ids
    kshape, vshape = kv_cache_shapes(NB, bs, H, D, torch.half)
    g = torch.Generator(device=dev).manual_seed(1234 + (rank % npw))
    kv = []
    for _ in range(L):
        if is_prompt:
            k = torch.randint(-2 ** 15, 2 ** 15 - 1, kshape, dtype=torch.int16, device=dev, generator=g).view(torch.half)
            v = torch.randint(-2 ** 15, 2 ** 15 - 1, vshape, dtype=torch.int16, device=dev, generator=g).view(torch.half)
        else:
            k = torch.zeros(kshape, dtype=torch.half, device=dev)
            v = torch.zeros(vshape, dtype=torch.half, device=dev)
        kv.append((k, v))
    block_bytes = H * D * bs * 2

    mgr = KVCacheCommManager(rank, world, npw, group=boot, device=dev)
    mgr.setup_comm(L, kv, use_comm_stream=comm_stream)
    comm = mgr.kvcache_comm
    assert comm.block_size == block_bytes
    comm.max_ctas = args.max_ctas
    lib = comm._lib
    spin_cycles = int(pace_us * 1e-6 * 1.9e9)

    def block_lists(pattern, seed):
        rnd = random.Random(seed)
        if pattern == "contiguous":
            return [list(range(p * blocks_per_prompt, (p + 1) * blocks_per_prompt)) for p in range(args.prompts)]
        ids = list(range(0, NB, 2))  # no two ids adjacent -> every run is a single block
        rnd.shuffle(ids)
        return [ids[p * blocks_per_prompt:(p + 1) * blocks_per_prompt] for p in range(args.prompts)]

    def checksum(blocks):
        """Sum over layers/K/V of the int64 words of the given blocks (wrapping) -- bit-exact fingerprint."""
        idx = torch.tensor(sorted(blocks), device=dev)
        tot = torch.zeros((), dtype=torch.int64, device=dev)
        for k, v in kv:
            tot += k.view(torch.int64).view(NB, -1)[idx].sum() + v.view(torch.int64).view(NB, -1)[idx].sum()
        return tot

    def total_sum():
        tot = torch.zeros((), dtype=torch.int64, device=dev)
        for k, v in kv:
            tot += k.view(torch.int64).sum() + v.view(torch.int64).sum()
        return tot

    def clear_token_side():
        if not is_prompt:
            for k, v in kv:
                k.zero_()
                v.zero_()

    def run_ref_api(lists):
        if is_prompt:
            for layer in range(L):
                for p, blocks in enumerate(lists):
                    for s, n in runs_of[id(blocks)]:
                        mgr.put(p % 256, layer, s, n)
            for p in range(len(lists)):
                mgr.signal_and_flush(p % 256)
        else:
            for p in range(len(lists)):
                mgr.wait(p % 256)

    def run_ours(lists, mode):
        """Returns the event that marks the end of the prompt side's compute (paced mode) or None."""
        if mode == "ref_api":
            return run_ref_api(lists)
        if mode == "paced":
            if is_prompt:
                for layer in range(L):
                    torch.cuda._sleep(spin_cycles)          # this layer's prefill compute (all prompts of the batch)
                    for p, blocks in enumerate(lists):     # the backends' loop (flash_attn.py:81-86): deferred puts,
                        for s, n in runs_of[id(blocks)]:   # one launch per layer for the whole batch
                            mgr.put(p % 256, layer, s, n)
                    mgr.flush()
                done = torch.cuda.Event(enable_timing=True)
                done.record()
                for p in range(len(lists)):
                    mgr.signal_and_flush(p % 256)
                return done
            for p in range(len(lists)):
                mgr.wait(p % 256)
            return None
        for p, blocks in enumerate(lists):
            sem = p % 256
            runs = runs_of[id(blocks)]
            if is_prompt:
                if mode == "bulk":
                    mgr.put_layers(sem, 0, L, runs)
                else:
                    for layer in range(L):
                        mgr.put_runs(sem, layer, runs)
                mgr.signal_and_flush(sem)
            else:
                mgr.wait(sem)
        return None

    def run_nccl(lists):
        works = []
        for blocks in lists:
            runs = runs_of[id(blocks)]
            ops = []
            for k, v in kv:
                for t in (k, v):
                    for s, n in runs:
                        ops.append(dist.P2POp(dist.isend if is_prompt else dist.irecv, t[s:s + n], peer))
            works += dist.batch_isend_irecv(ops)
        for w in works:
            w.wait()

    results = []
    configs = list(sweep) or [(-1, args.max_ctas, comm.layers_per_launch)]
    for pattern in args.patterns.split(","):
        for mode in args.modes.split(","):
            for variant, ctas, lpl in (configs if mode != "nccl" else configs[:1]):
                if mode != "nccl":
                    lib.b200_kv_put_configure(variant, 0)
                    comm.max_ctas = ctas
                    comm.layers_per_launch = max(1, lpl)
                lists = block_lists(pattern, 7)
                if mode == "nccl" and pattern == "fragmented":
                    lists = lists[:args.nccl_max_prompts]
                # the scheduler coalesces a prompt's block list once (vllm/utils.py:403-426); not part of the transfer
                runs_of = {id(b): tuple(map(tuple, coalesce_blocks(b))) for b in lists}
                nbytes = sum(len(b) for b in lists) * block_bytes * 2 * L
                times, tails = [], []
                ok = True
                for it in range(args.warmup + args.iters):
                    clear_token_side()
                    torch.cuda.synchronize()
                    dist.barrier()
                    beg, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    beg.record()
                    done = None
                    if mode == "nccl":
                        run_nccl(lists)
                    else:
                        done = run_ours(lists, mode)
                        comm.join_current_stream()   # side stream: the end event must see the puts and signals
                    end.record()
                    torch.cuda.synchronize()
                    ms = torch.tensor([beg.elapsed_time(end), done.elapsed_time(end) if done is not None else 0.0],
                                      dtype=torch.float64, device=dev)
                    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
                    if it >= args.warmup:
                        times.append(ms[0].item())
                        tails.append(ms[1].item())
                    # ---- verify: moved blocks match the sender's fingerprint, nothing else was touched ----
                    moved = [b for bl in lists for b in bl]
                    fp = checksum(moved)
                    if is_prompt:
                        dist.send(fp.view(1), peer)
                    else:
                        theirs = torch.zeros(1, dtype=torch.int64, device=dev)
                        dist.recv(theirs, peer)
                        ok = ok and bool(theirs.item() == fp.item()) and bool(total_sum().item() == fp.item())
                        if mode != "nccl":
                            ok = ok and not comm.timed_out()
                okt = torch.tensor([1 if ok else 0], device=dev)
                dist.all_reduce(okt, op=dist.ReduceOp.MIN)
                best, mean = min(times), sum(times) / len(times)
                per_pair = nbytes / (mean * 1e-3) / 1e9
                line = {"metric": "kv_transfer_bandwidth_per_gpu_pair", "value": round(per_pair, 2), "unit": "GB/s",
                        "n_gpus": world, "pairs": npw, "pattern": pattern, "mode": mode, "prompts_in_flight": len(lists),
                        "bytes_per_pair": nbytes, "ms_mean": round(mean, 4), "ms_best": round(best, 4),
                        "best_GBps": round(nbytes / (best * 1e-3) / 1e9, 2),
                        "aggregate_GBps": round(per_pair * npw, 2),
                        "frac_of_nvlink_measured_770": round(per_pair / NVLINK_MEASURED_GBS, 4),
                        "frac_of_nvlink_nominal_900": round(per_pair / NVLINK_NOMINAL_GBS, 4),
                        "runs_per_prompt": len(coalesce_blocks(lists[0])), "bit_exact": bool(okt.item()),
                        "config": {"workload": f"C5 Llama-2-13B tp=4 shard: {L} layers, {H} kv heads, head {D}, block {bs}, fp16, "
                                               f"{args.prompt_tokens}-token prompts, {npw} prompt -> {npw} token GPUs"}}
                if mode != "nccl":
                    import ctypes
                    cv, cc = ctypes.c_int32(), ctypes.c_int32()
                    lib.b200_kv_put_config(ctypes.byref(cv), ctypes.byref(cc))   # what the library actually used
                    line.update({"put_variant": {0: "simt", 1: "bulk"}[cv.value],
                                 "max_ctas": ctas if ctas > 0 else cc.value,
                                 "layers_per_launch": comm.layers_per_launch, "comm_stream": comm.comm_stream is not None})
                if mode == "paced":
                    tail = sum(tails) / len(tails)
                    line.update({"time_to_signal_ms": round(tail, 4), "pace_us_per_layer": pace_us,
                                 "note": "value = bytes / whole round (dominated by the stand-in compute); "
                                         "time_to_signal_ms = end of the last layer's compute -> signal delivered "
                                         f"(one layer of all prompts = {nbytes / L / 1e6:.1f} MB)"})
                results.append(line)
                if rank == 0 and echo:
                    print(json.dumps(line), flush=True)
    dist.barrier()
    mgr.destroy_comm()
    return results


if __name__ == "__main__":
    sys.exit(main())
