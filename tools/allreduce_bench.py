"""One-shot NVLink all-reduce under torchrun, one process per GPU: bit-exact check against the oracle
and latency against NCCL all_reduce for the decode activation sizes [batch, hidden] of the BASELINE configs.
    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/allreduce_bench.py
"""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import paged_kv_oracle as O  # noqa: E402  (checker only)
from vllm_splitwise_b200.custom_all_reduce import CustomAllreduce  # noqa: E402


CASES = (("c2_decode_64x4096_f16", 64 * 4096, torch.float16),
         ("c3_decode_256x8192_bf16", 256 * 8192, torch.bfloat16),
         ("single_token_1x8192_bf16", 8192, torch.bfloat16),
         ("prefill_2048x5120_f16", 2048 * 5120 // 2, torch.float16))


def run(rank, world, dev, group, iters=200, echo=False):
    """Every case of CASES on an initialised NCCL process group; returns the result lines (bench.py's
    `extra.allreduce`)."""
    car = CustomAllreduce(rank, world, dev, group=group, max_size=16 << 20)
    lines = []
    for name, numel, dtype in CASES:
        xs = [(torch.rand(numel, generator=torch.Generator().manual_seed(r)) * 4 - 2).to(dtype) for r in range(world)]
        x = xs[rank].to(dev)
        out = car.all_reduce(x)
        torch.cuda.synchronize()
        exact = bool(torch.equal(out.cpu().view(torch.int16 if dtype != torch.float32 else torch.int32),
                                 O.all_reduce_sum(xs).view(torch.int16 if dtype != torch.float32 else torch.int32)))

        def timed(fn, iters=iters):
            for _ in range(5):
                fn()
            dist.barrier(); torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(iters):
                fn()
            e1.record(); torch.cuda.synchronize()
            t = torch.tensor([e0.elapsed_time(e1) / iters * 1000], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return t.item()

        ours = timed(lambda: car.all_reduce(x, out))
        xr = car.registered_input(x.shape, dtype)
        xr.copy_(x)
        ours_reg = timed(lambda: car.all_reduce(xr, out))
        torch.cuda.synchronize()
        exact = exact and bool(torch.equal(out.cpu().view(torch.int16), O.all_reduce_sum(xs).view(torch.int16)))
        y = x.clone()
        nccl = timed(lambda: dist.all_reduce(y))
        flags = torch.tensor([int(exact), int(not car.timed_out())], device=dev)
        dist.all_reduce(flags, op=dist.ReduceOp.MIN)
        nbytes = numel * x.element_size()
        line = {"case": name, "n_gpus": world, "bytes": nbytes, "oneshot_us": round(ours, 2),
                "oneshot_registered_input_us": round(ours_reg, 2), "nccl_us": round(nccl, 2),
                "speedup_vs_nccl": round(nccl / ours, 2), "registered_speedup_vs_nccl": round(nccl / ours_reg, 2),
                "remote_read_GBps_per_gpu": round(nbytes * (world - 1) / ours / 1e3, 1),
                "bit_exact_all_ranks": bool(flags[0].item()), "timed_out": not bool(flags[1].item()),
                "includes": "copy into the registered buffer + one kernel (start flags, sum, end flags)"}
        lines.append(line)
        if rank == 0 and echo:
            print(json.dumps(line), flush=True)
    car.close()
    return lines


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dev = torch.device("cuda", torch.cuda.current_device())
    dist.init_process_group("nccl", device_id=dev)
    run(rank, world, dev, dist.group.WORLD, echo=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
