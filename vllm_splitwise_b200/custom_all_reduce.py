"""One-shot all-reduce over NVLink peer memory: host-side mirror of the reference's
CustomAllreduce (vllm/model_executor/parallel_utils/custom_all_reduce.py:149-233): `should_custom_ar`, `all_reduce`
(the unregistered-buffer path: copy into the registered buffer, then reduce), `close`.  One process per GPU; handles
travel once over the worker's torch.distributed group (all_gather_object), like the reference's
`_gather_ipc_meta` (:192-204).  No NCCL on the data path, no CPU fallback.
"""
import ctypes
from typing import List, Optional, Sequence

import torch

from . import _lib
from ._C._common import dtype_code, stream_ptr


_MAX_CTAS = 160  # rows of the per-CTA barrier array (kMaxCtas2 in csrc/all_reduce.cu)


class AllReduceTimeout(RuntimeError):
    pass


class CustomAllreduce:

    def __init__(self, rank: int, world_size: int, device, group=None, max_size: int = 8 << 20,
                 peers: Optional[Sequence["CustomAllreduce"]] = None) -> None:
        """`group`: torch.distributed group used once to exchange IPC handles (one process per GPU).
        `peers`: instead, communicators living in THIS process (one per GPU, peer access) -- tests and loopback."""
        self.rank, self.world_size = rank, world_size
        self.device = torch.device(device)
        self.max_size = max_size
        # registered input (first half) and the two-stage kernel's scratch (second half), one allocation
        self.buffer = torch.zeros(2 * max_size, dtype=torch.uint8, device=self.device)
        # start[world], done[world], mid[160][world]
        self.flags = torch.zeros(2 + _MAX_CTAS, world_size, dtype=torch.int64, device=self.device)
        # messages of at least this many bytes go through reduce-scatter + all-gather when world_size > 2
        # (the reference switches at 512 KiB / 256 KiB for 4 / 8 ranks, custom_all_reduce.cuh:452-463)
        self.two_stage_min_bytes = 1 << 20
        self.counter = torch.zeros(1, dtype=torch.int32, device=self.device)
        self.timeout = torch.zeros(1, dtype=torch.int32, device=self.device)
        self.epoch = 0
        self.max_ctas = 0
        self._opened: List[int] = []
        self._in = (ctypes.c_int64 * world_size)()
        self._start = (ctypes.c_int64 * world_size)()
        self._done = (ctypes.c_int64 * world_size)()
        self._tmp = (ctypes.c_int64 * world_size)()
        self._mid = (ctypes.c_int64 * world_size)()
        if group is not None:
            self._connect_ipc(group)

    # -- wiring ------------------------------------------------------------------------------
    def _fill(self, r: int, buf_ptr: int, flags_ptr: int) -> None:
        self._in[r] = buf_ptr
        self._start[r] = flags_ptr + 8 * self.rank                      # start_of_rank_r[my rank]
        self._done[r] = flags_ptr + 8 * (self.world_size + self.rank)   # done_of_rank_r[my rank]
        self._tmp[r] = buf_ptr + self.max_size
        self._mid[r] = flags_ptr + 8 * (2 * self.world_size + self.rank)  # mid_of_rank_r[CTA 0][my rank]

    @staticmethod
    def connect_local(comms: Sequence["CustomAllreduce"]) -> None:
        """All ranks in one process: plain device pointers (peer access must be enabled between the devices)."""
        lib = _lib.load()
        for a in comms:
            for b in comms:
                if a.device != b.device:
                    _lib.check(lib.b200_kv_enable_peer_access(a.device.index, b.device.index), "enable peer access")
                a._fill(b.rank, b.buffer.data_ptr(), b.flags.data_ptr())

    def _export(self, t: torch.Tensor):
        handle = (ctypes.c_ubyte * _lib.IPC_HANDLE_BYTES)()
        off, size = ctypes.c_int64(), ctypes.c_int64()
        _lib.check(_lib.load().b200_kv_ipc_export(t.data_ptr(), handle, ctypes.byref(off), ctypes.byref(size)),
                   "ipc export")
        return bytes(handle), off.value

    def _connect_ipc(self, group) -> None:
        import torch.distributed as dist
        mine = {"rank": self.rank, "buf": self._export(self.buffer), "flags": self._export(self.flags)}
        infos = [None] * self.world_size
        dist.all_gather_object(infos, mine, group=group)
        cache = {}

        def opened(handle: bytes) -> int:
            if handle not in cache:
                buf = (ctypes.c_ubyte * _lib.IPC_HANDLE_BYTES).from_buffer_copy(handle)
                base = ctypes.c_void_p()
                _lib.check(_lib.load().b200_kv_ipc_open(buf, ctypes.byref(base)), "ipc open")
                cache[handle] = base.value
                self._opened.append(base.value)
            return cache[handle]

        for info in infos:
            r = info["rank"]
            if r == self.rank:
                self._fill(r, self.buffer.data_ptr(), self.flags.data_ptr())
            else:
                (bh, boff), (fh, foff) = info["buf"], info["flags"]
                self._fill(r, opened(bh) + boff, opened(fh) + foff)
        dist.barrier(group=group)

    # -- the op --------------------------------------------------------------------------------
    def should_custom_ar(self, inp: torch.Tensor) -> bool:
        """reference custom_all_reduce.py:101-117 (size and alignment rule; the one-shot kernel serves every size here)"""
        nbytes = inp.numel() * inp.element_size()
        return inp.is_cuda and inp.is_contiguous() and nbytes % 16 == 0 and nbytes <= self.max_size

    def all_reduce(self, inp: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        if not self.should_custom_ar(inp):
            raise RuntimeError("all_reduce: tensor must be a contiguous CUDA tensor of a multiple of 16 bytes, at most "
                               f"{self.max_size} bytes")
        if out is None:
            out = torch.empty_like(inp)
        nbytes = inp.numel() * inp.element_size()
        if inp.data_ptr() != self.buffer.data_ptr():   # registered input (all_reduce_reg, :218-224): no staging copy
            self.buffer[:nbytes].copy_(inp.view(-1).view(torch.uint8), non_blocking=True)  # all_reduce_unreg (:226-229)
        self.epoch += 1
        if self.world_size > 2 and nbytes >= self.two_stage_min_bytes:
            st = _lib.load().b200_allreduce_twostage(
                self._in, self._tmp, self._start, self._done, self._mid, self.flags.data_ptr(),
                self.flags.data_ptr() + 8 * self.world_size, self.flags.data_ptr() + 16 * self.world_size,
                self.counter.data_ptr(), self.timeout.data_ptr(), out.data_ptr(), inp.numel(), dtype_code(inp, "input"),
                self.rank, self.world_size, self.epoch, 0, stream_ptr(self.device))
            _lib.check(st, "all_reduce (two-stage)")
            return out
        st = _lib.load().b200_allreduce_oneshot(
            self._in, self._start, self._done, self.flags.data_ptr(), self.flags.data_ptr() + 8 * self.world_size,
            self.counter.data_ptr(), self.timeout.data_ptr(), out.data_ptr(), inp.numel(), dtype_code(inp, "input"),
            self.rank, self.world_size, self.epoch, 0, self.max_ctas, stream_ptr(self.device))
        _lib.check(st, "all_reduce")
        return out

    def registered_input(self, shape, dtype: torch.dtype) -> torch.Tensor:
        """A view of the registered buffer: a producer (e.g. the o_proj GEMM) that writes its result here can call
        all_reduce on it without the staging copy -- the reference's registered-buffer path (:206-224)."""
        n = 1
        for d in shape:
            n *= d
        nbytes = n * torch.tensor([], dtype=dtype).element_size()
        if nbytes > self.max_size:
            raise RuntimeError("registered_input: larger than the registered buffer")
        return self.buffer[:nbytes].view(dtype).view(*shape)  # the first half of the allocation is the input

    def timed_out(self) -> bool:
        return bool(self.timeout.item())

    def check_timeout(self) -> None:
        """Call at a step boundary (synchronises).  A kernel whose bounded flag spin expired completed WITHOUT all
        peers' data, so the sums of that call are wrong: raise, after clearing the flag.  The protocol itself
        needs no repair -- epochs are monotonic, a late peer's flag satisfies every later call."""
        if self.timed_out():
            self.timeout.zero_()
            raise AllReduceTimeout(f"custom all-reduce on rank {self.rank}: a peer did not arrive within the spin "
                                   "bound; the results of the affected call(s) are invalid")

    def close(self) -> None:
        for base in self._opened:
            _lib.load().b200_kv_ipc_close(base)
        self._opened = []
