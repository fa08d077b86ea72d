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


def load_reference_custom_ar():
    """The UNMODIFIED reference all-reduce (csrc/custom_all_reduce.cu/.cuh) from oracle/_ref, or None if not built."""
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
    from _util import load_ref_cuda
    ref = load_ref_cuda()
    return getattr(ref, "custom_ar", None) if ref is not None else None


class ReferenceAllreduce:
    """Drives the reference kernels the way vllm/model_executor/parallel_utils/custom_all_reduce.py:146-232 does:
    a meta buffer (signals + scratch) and a registered staging buffer, IPC handles exchanged through the group."""

    def __init__(self, ar, rank, world, dev, group, max_size):
        self.ar, self.group, self.world = ar, group, world
        self.meta = torch.zeros(ar.meta_size() + max_size, dtype=torch.uint8, device=dev)
        self.buffer = torch.empty(max_size, dtype=torch.uint8, device=dev)
        self.rank_data = torch.empty(8 << 20, dtype=torch.uint8, device=dev)
        handles, offsets = self._ipc_meta(self.meta)
        self.ptr = ar.init_custom_ar(self.meta, self.rank_data, handles, offsets, rank, True)
        handles, offsets = self._ipc_meta(self.buffer)
        ar.register_buffer(self.ptr, self.buffer, handles, offsets)

    def _ipc_meta(self, t):
        # (handle of the allocation's base, offset of t in it) -- cudaIpcGetMemHandle through the library's own export
        # helper; torch's storage._share_cuda_() handle is not a bare cudaIpcMemHandle_t any more (the reference's
        # cudaIpcOpenMemHandle rejects it with "invalid argument" on torch 2.11).
        import ctypes
        from vllm_splitwise_b200 import _lib
        handle = (ctypes.c_ubyte * _lib.IPC_HANDLE_BYTES)()
        off, size = ctypes.c_int64(), ctypes.c_int64()
        _lib.check(_lib.load().b200_kv_ipc_export(t.data_ptr(), handle, ctypes.byref(off), ctypes.byref(size)), "ipc export")
        everyone = [None] * self.world
        dist.all_gather_object(everyone, (bytes(handle), off.value), group=self.group)
        return [e[0] for e in everyone], [e[1] for e in everyone]

    def all_reduce(self, x, out):
        self.ar.all_reduce_unreg(self.ptr, x, self.buffer, out)
        return out

    def close(self):
        self.ar.dispose(self.ptr)


CASES = (("c2_decode_64x4096_f16", 64 * 4096, torch.float16),
         ("c3_decode_256x8192_bf16", 256 * 8192, torch.bfloat16),
         ("single_token_1x8192_bf16", 8192, torch.bfloat16),
         ("prefill_2048x5120_f16", 2048 * 5120 // 2, torch.float16))


def run(rank, world, dev, group, iters=200, echo=False):
    """Every case of CASES on an initialised NCCL process group; returns the result lines (bench.py's
    `extra.allreduce`)."""
    car = CustomAllreduce(rank, world, dev, group=group, max_size=16 << 20)
    ref_mod = load_reference_custom_ar()
    have_ref = torch.tensor([1 if ref_mod is not None else 0], device=dev)
    dist.all_reduce(have_ref, op=dist.ReduceOp.MIN)
    gloo = dist.new_group(backend="gloo")   # the IPC handles are pickled objects
    ref = ReferenceAllreduce(ref_mod, rank, world, dev, gloo, 16 << 20) if have_ref.item() and world % 2 == 0 else None
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
        ref_us, same_as_ref = None, None
        if ref is not None:
            ref_out = torch.empty_like(x)
            ref.all_reduce(x, ref_out)
            torch.cuda.synchronize()
            same = torch.tensor([int(torch.equal(ref_out.view(torch.int16), out.view(torch.int16)))], device=dev)
            dist.all_reduce(same, op=dist.ReduceOp.MIN)
            same_as_ref = bool(same.item())
            ref_us = timed(lambda: ref.all_reduce(x, ref_out))
        flags = torch.tensor([int(exact), int(not car.timed_out())], device=dev)
        dist.all_reduce(flags, op=dist.ReduceOp.MIN)
        nbytes = numel * x.element_size()
        line = {"case": name, "n_gpus": world, "bytes": nbytes, "oneshot_us": round(ours, 2),
                "oneshot_registered_input_us": round(ours_reg, 2), "nccl_us": round(nccl, 2),
                "speedup_vs_nccl": round(nccl / ours, 2), "registered_speedup_vs_nccl": round(nccl / ours_reg, 2),
                "remote_read_GBps_per_gpu": round(nbytes * (world - 1) / ours / 1e3, 1),
                "bit_exact_all_ranks": bool(flags[0].item()), "timed_out": not bool(flags[1].item()),
                "reference_cuda_us": round(ref_us, 2) if ref_us is not None else None,
                "speedup_vs_reference_cuda": round(ref_us / ours, 2) if ref_us else None,
                "bit_identical_to_reference_cuda": same_as_ref,
                "includes": "copy into the registered buffer + one kernel (start flags, sum, end flags)"}
        lines.append(line)
        if rank == 0 and echo:
            print(json.dumps(line), flush=True)
    if ref is not None:
        ref.close()
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
