"""Splitwise prefill -> decode KV-block transfer: mirror of the reference's
vllm/worker/comm_utils.py (KVCacheCommunicator :93-147, KVCacheCommManager :150-222) over NVLink
peer memory instead of MSCCL++ proxy channels.

Same call surface -- `KVCacheCommManager(rank, world_size, num_prompt_workers, init_method)`,
`.setup_comm(num_layers, kv_cache)`, `.put(sem_id, layer_id, block_start, num_blocks)`,
`.signal_and_flush(sem_id)`, `.wait(sem_id)`, `.destroy_comm()` -- and the same meaning: after
`wait(sem)` completes on the token GPU's stream, every byte `put` under `sem` since the previous
signal is visible in the token GPU's caches at the same block offsets.

What differs is the mechanism.  The reference launches a 1-thread kernel per put that posts a
trigger to a host FIFO; a host proxy thread turns it into a copy (csrc/kv_comm_kernels.cu:11-20).
Here the token GPU's cache tensors are mapped into the prompt process once (CUDA IPC, or plain
peer access inside one process) and `put` is ONE multi-CTA kernel that stores the block runs
straight into the peer mapping; `signal` is a system-scope release store to a flag in the token
GPU's memory and `wait` an acquire spin on it.  `put_runs` / `put_layers` batch any number of
runs and layers into a single launch (the reference needs one launch per run per K/V).
"""
import ctypes
from typing import List, Optional, Sequence, Tuple

import torch

from . import _lib
from .utils import MAX_SLOT_IDS

HEAD_TYPES = [0, 1]  # 0 = keys, 1 = values (reference comm_utils.py:16)
PUT_CHAINED = 1 << 30  # include/b200_kv.h B200_KV_PUT_CHAINED


def pair_rank(rank: int, world_size: int, num_prompt_workers: int) -> int:
    """Prompt rank r is paired with token rank (r + num_prompt_workers) % world (comm_utils.py:165)."""
    return (rank + num_prompt_workers) % world_size


def is_prompt_rank(rank: int, num_prompt_workers: int) -> bool:
    return rank < num_prompt_workers


def exchange_peer_info(group, rank: int, world_size: int, num_prompt_workers: int, mine: dict) -> Optional[dict]:
    """All ranks publish `mine` (IPC handles of their cache buffers and flags); a prompt rank gets back
    the record of its token rank (pair_rank), a token rank gets None.  Collective over `group` (any
    torch.distributed backend; the records are pickled, so gloo is enough).  This replaces the MSCCL++
    bootstrap + `make_connection` + `register_memory` exchange (reference comm_utils.py:158-197)."""
    import torch.distributed as dist
    everyone: List[Optional[dict]] = [None] * world_size
    dist.all_gather_object(everyone, mine, group=group)
    for r, rec in enumerate(everyone):
        if rec is None or rec.get("rank") != r:
            raise RuntimeError(f"KV comm bootstrap: bad record from rank {r}")
    if not is_prompt_rank(rank, num_prompt_workers):
        return None
    return everyone[pair_rank(rank, world_size, num_prompt_workers)]


_runs_cache: dict = {}


def _runs_array(runs: Sequence[Tuple[int, int]]):
    """ctypes int32 [num_runs][2] for the C ABI.  The worker puts the same coalesced block list once per
    layer (reference flash_attn.py:81-86), so the last few conversions are cached by content."""
    key = tuple(map(tuple, runs)) if not isinstance(runs, tuple) else runs
    hit = _runs_cache.get(key)
    if hit is not None:
        return hit
    flat: List[int] = []
    for start, n in runs:
        flat += (int(start), int(n))
    out = ((ctypes.c_int32 * len(flat))(*flat), len(flat) // 2)
    if len(_runs_cache) >= 64:
        _runs_cache.pop(next(iter(_runs_cache)))
    _runs_cache[key] = out
    return out


class KVWaitTimeout(RuntimeError):
    """A KV-transfer wait() ran out of polls: the prompt side never signalled these semaphores."""

    def __init__(self, sem_ids) -> None:
        super().__init__(f"KV wait timed out on semaphore(s) {list(sem_ids)}: the blocks of these sequences did "
                         "not arrive; results computed from them are invalid")
        self.sem_ids = list(sem_ids)


class KVCacheCommunicator:
    """put / signal / wait on one prompt<->token pair.

    block_size: bytes of one block of one layer's K (== V), as in the reference (comm_utils.py:199-201).
    src_table / dst_table: device int64 [2*num_layers] base addresses, order [L0.K, L0.V, L1.K, ...];
    dst entries are addresses of the PEER's tensors as mapped into this process.
    remote_flags: address (in this process) of the peer's uint64[MAX_SLOT_IDS] flag array.
    local_flags: this GPU's flag array (what `wait` polls).
    """

    def __init__(self, block_size: int, num_layers: int, device: torch.device,
                 src_table: Optional[torch.Tensor], dst_table: Optional[torch.Tensor],
                 remote_flags: Optional[int], local_flags: torch.Tensor, my_rank: int, remote_rank: int,
                 max_ctas: int = 0, wait_max_spins: int = 100000000, defer_puts: bool = True,
                 comm_stream: Optional["torch.cuda.Stream"] = None, layers_per_launch: int = 1) -> None:
        self.block_size = block_size
        self.num_layers = num_layers
        self.device = device
        self.src_table = src_table
        self.dst_table = dst_table
        self.remote_flags = remote_flags
        self.local_flags = local_flags
        self.my_rank = my_rank
        self.remote_rank = remote_rank
        self.max_ctas = max_ctas
        self.wait_max_spins = wait_max_spins
        self.sent = [0] * MAX_SLOT_IDS       # signals issued per semaphore (prompt side)
        self.expected = [0] * MAX_SLOT_IDS   # signals awaited per semaphore (token side)
        # one sticky flag per semaphore, set by a wait kernel whose bounded spin expired (csrc/kv_transfer.cu);
        # read by check_timeouts() / timed_out()
        self.timeout_flags = torch.zeros(MAX_SLOT_IDS, dtype=torch.int32, device=device)
        self._lib = _lib.load()
        # put() batching: the attention backends call put() once per coalesced run per semaphore per layer
        # (reference flash_attn.py:81-86).  With defer_puts the runs of one layer are collected on the host
        # and leave as ONE launch when the layer changes, on flush(), or before signal/wait -- always on the
        # same stream and before the signal, so the reference's ordering contract is unchanged.
        self.defer_puts = defer_puts
        self._pending_layer = -1
        self._pending_runs: List[Tuple[int, int]] = []
        # layers_per_launch > 1: consecutive layers that put the SAME run list (what the backends do: one block
        # list per batch, every layer) are merged into one launch of up to that many layers.  A launch then moves
        # tens of MB instead of ~10 MB, so neither the host's launch rate nor the gaps between kernels limit the
        # link; the price is that a layer's blocks leave up to layers_per_launch - 1 layers later.
        self.layers_per_launch = max(1, int(layers_per_launch))
        self._batch_first, self._batch_count, self._batch_runs = -1, 0, ()
        self._tables: dict = {}
        self._src_base = src_table.data_ptr() if src_table is not None else 0
        self._dst_base = dst_table.data_ptr() if dst_table is not None else 0
        self._dev_index = device.index if device.index is not None else torch.cuda.current_device()
        self._join = torch.cuda.Event() if comm_stream is not None else None
        # optional side stream: transfers then overlap the prefill compute that follows on the main stream
        # (the reference's host proxy copies asynchronously too); signal rides the same side stream.
        self.comm_stream = comm_stream
        self._fence = torch.cuda.Event() if comm_stream is not None else None

    def _stream(self) -> int:
        if self.comm_stream is not None:
            return self.comm_stream.cuda_stream
        return torch.cuda.current_stream(self.device).cuda_stream

    def _order_after_current(self) -> None:
        """Side stream only: what the caller's stream has enqueued so far (the cache writes) happens first."""
        if self.comm_stream is not None:
            self._fence.record(torch.cuda.current_stream(self.device))
            self.comm_stream.wait_event(self._fence)

    def join_current_stream(self) -> None:
        """Side stream only: the caller's current stream waits for everything issued to the comm stream so far."""
        if self.comm_stream is not None:
            self._join.record(self.comm_stream)
            torch.cuda.current_stream(self.device).wait_event(self._join)

    def _close_layer(self) -> None:
        """The runs collected for the pending layer are complete: merge them into the batch of consecutive
        layers with the same run list, launching the batch when it is full or cannot be extended."""
        if not self._pending_runs:
            return
        layer, runs = self._pending_layer, tuple(self._pending_runs)
        self._pending_layer, self._pending_runs = -1, []
        self._add_layer(layer, runs)

    def _add_layer(self, layer: int, runs: tuple) -> None:
        if self._batch_count and layer == self._batch_first + self._batch_count and \
                (runs is self._batch_runs or runs == self._batch_runs):
            self._batch_count += 1
        else:
            self._launch_batch()
            self._batch_first, self._batch_count, self._batch_runs = layer, 1, runs
        if self._batch_count >= self.layers_per_launch:
            self._launch_batch()

    def _launch_batch(self) -> None:
        if self._batch_count:
            first, count, runs = self._batch_first, self._batch_count, self._batch_runs
            self._batch_count = 0
            self._order_after_current()
            self._put(2 * first, 2 * count, runs)

    def flush(self) -> None:
        """Launch the deferred puts (the pending layer's runs and any batch of merged layers)."""
        self._close_layer()
        self._launch_batch()

    def _check_sem(self, sem_id: int) -> None:
        if not 0 <= sem_id < MAX_SLOT_IDS:
            raise ValueError(f"sem_id {sem_id} out of range [0, {MAX_SLOT_IDS})")

    PARAM_TABLE_RUNS = 440  # runs one launch carries in its kernel parameters (csrc/kv_transfer.cu kMaxRuns)

    def _device_table(self, key):
        """(device int32 table, num_runs, tile_vecs, tiles_per_buf) for a long run list, uploaded once and kept:
        the same list is put for every layer of the batch."""
        hit = self._tables.get(key)
        if hit is not None:
            return hit
        arr, n = _runs_array(key)
        host = torch.empty(3 * n + 1, dtype=torch.int32).pin_memory()
        tv, tpb = ctypes.c_int32(), ctypes.c_int64()
        _lib.check(self._lib.b200_kv_put_table_build(arr, n, self.block_size, host.data_ptr(), ctypes.byref(tv),
                                                     ctypes.byref(tpb)), "kv put table")
        with torch.cuda.device(self.device):
            stream = self.comm_stream if self.comm_stream is not None else torch.cuda.current_stream(self.device)
            with torch.cuda.stream(stream):
                dev = host.to(self.device, non_blocking=True)
        if len(self._tables) >= 16:
            self._tables.pop(next(iter(self._tables)))
        out = (dev, n, tv.value, tpb.value, host)   # host kept alive until the async copy is long done
        self._tables[key] = out
        return out

    def _put(self, first_buf: int, num_bufs: int, runs: Sequence[Tuple[int, int]]) -> None:
        if self.src_table is None or self.dst_table is None:
            raise RuntimeError("put() called on a rank that has no outgoing KV connection")
        key = runs if isinstance(runs, tuple) else tuple(map(tuple, runs))
        if len(key) == 0:
            return
        if self.comm_stream is not None:
            # Side stream: every put is ordered behind the caller's cache writes by an event (_order_after_current),
            # not by stream order behind the previous put, so consecutive puts may overlap (B200_KV_PUT_CHAINED).
            stream, ctas = self.comm_stream.cuda_stream, max(0, self.max_ctas) | PUT_CHAINED
        else:
            stream, ctas = torch.cuda.current_stream(self.device).cuda_stream, self.max_ctas
        # this is the per-layer hot call of a prefill worker: switch devices only when the caller sits on another one
        switch = torch.cuda.current_device() != self._dev_index
        if switch:
            prev = torch.cuda.current_device()
            torch.cuda.set_device(self._dev_index)
        try:
            if len(key) > self.PARAM_TABLE_RUNS:
                dev, n, tv, tpb, _ = self._device_table(key)
                st = self._lib.b200_kv_put_blocks_table(self._src_base + 8 * first_buf, self._dst_base + 8 * first_buf,
                                                        num_bufs, dev.data_ptr(), n, tv, tpb, self.block_size,
                                                        ctas, stream)
            else:
                arr, n = _runs_array(key)
                st = self._lib.b200_kv_put_blocks(self._src_base + 8 * first_buf, self._dst_base + 8 * first_buf,
                                                  num_bufs, arr, n, self.block_size, ctas, stream)
        finally:
            if switch:
                torch.cuda.set_device(prev)
        _lib.check(st, "kv put")

    # ---- reference surface -------------------------------------------------------------
    def put(self, sem_id: int, layer_id: int, block_start: int, num_blocks: int) -> None:
        """K and V blocks [block_start, block_start+num_blocks) of one layer (comm_utils.py:132-147).
        Called once per coalesced run per semaphore per layer by the backends, i.e. tens of thousands of times
        per prefill batch: the deferred path is an append; everything else happens once per layer in flush()."""
        if layer_id != self._pending_layer or not self.defer_puts:
            self._put_slow(sem_id, layer_id, block_start, num_blocks)
        elif num_blocks > 0 and 0 <= sem_id < MAX_SLOT_IDS:
            self._pending_runs.append((block_start, num_blocks))
        else:
            self._check_sem(sem_id)

    def _put_slow(self, sem_id: int, layer_id: int, block_start: int, num_blocks: int) -> None:
        self._check_sem(sem_id)
        if not 0 <= layer_id < self.num_layers:
            raise ValueError(f"layer_id {layer_id} out of range [0, {self.num_layers})")
        if self.src_table is None or self.dst_table is None:
            raise RuntimeError("put() called on a rank that has no outgoing KV connection")
        if num_blocks <= 0:
            return
        if not self.defer_puts:
            self._order_after_current()
            self._put(2 * layer_id, 2, ((int(block_start), int(num_blocks)),))
            return
        self._close_layer()
        self._pending_layer = layer_id
        self._pending_runs.append((block_start, num_blocks))

    def signal_and_flush(self, sem_id: int) -> None:
        self._check_sem(sem_id)
        if self.remote_flags is None:
            raise RuntimeError("signal_and_flush() called on a rank that has no outgoing KV connection")
        self.flush()
        self._order_after_current()
        self.sent[sem_id] += 1
        with torch.cuda.device(self.device):
            st = self._lib.b200_kv_signal(self.remote_flags + 8 * sem_id, self.sent[sem_id], self._stream())
        _lib.check(st, "kv signal")

    def wait(self, sem_id: int) -> None:
        self._check_sem(sem_id)
        self.flush()
        self.expected[sem_id] += 1
        with torch.cuda.device(self.device):
            # always on the caller's stream: the decode kernels that follow on it must see the blocks
            st = self._lib.b200_kv_wait(self.local_flags.data_ptr() + 8 * sem_id, self.expected[sem_id],
                                        self.wait_max_spins, self.timeout_flags.data_ptr() + 4 * sem_id,
                                        torch.cuda.current_stream(self.device).cuda_stream)
        _lib.check(st, "kv wait")

    # ---- batched extensions ------------------------------------------------------------
    def put_runs(self, sem_id: int, layer_id: int, runs: Sequence[Tuple[int, int]]) -> None:
        """All coalesced runs of one layer (K and V) in one launch (merged with the neighbouring layers' identical
        run lists when layers_per_launch > 1; pass the same tuple object every layer to make that check free)."""
        self._check_sem(sem_id)
        if self._pending_runs:
            self._close_layer()
        self._add_layer(layer_id, runs if isinstance(runs, tuple) else tuple(map(tuple, runs)))

    def put_layers(self, sem_id: int, first_layer: int, num_layers: int, runs: Sequence[Tuple[int, int]]) -> None:
        """The same runs for a range of layers (K and V) in one launch."""
        self._check_sem(sem_id)
        self.flush()
        self._order_after_current()
        self._put(2 * first_layer, 2 * num_layers, runs)

    def timed_out(self) -> bool:
        """True if any wait() expired since the last check_timeouts()/reset (synchronises)."""
        return bool(self.timeout_flags.any().item())

    def reset_semaphore(self, sem_id: int) -> None:
        """Abort / resynchronise one semaphore after a missed or cancelled transfer (synchronises): the next
        wait(sem_id) expects the signal after the last one that actually ARRIVED, and the timeout flag is cleared.
        Without this a single missed signal would leave `expected` one ahead for ever and every later wait on the
        semaphore would spin out its full bound."""
        self._check_sem(sem_id)
        torch.cuda.current_stream(self.device).synchronize()
        self.expected[sem_id] = int(self.local_flags[sem_id].item())
        self.timeout_flags[sem_id] = 0

    def check_timeouts(self) -> None:
        """Call at a step boundary on the token side (synchronises).  If a wait() expired, the decode kernels
        enqueued behind it ran over blocks that never arrived: raise KVWaitTimeout naming the semaphores, after
        resynchronising them (reset_semaphore) so that the worker can drop those sequences and carry on.
        The reference's MSCCL++ wait(maxSpinCount) aborts the process on expiry (kv_comm_kernels.cu:7)."""
        flags = self.timeout_flags.cpu()
        bad = flags.nonzero().flatten().tolist()
        if not bad:
            return
        for sem in bad:
            self.reset_semaphore(sem)
        raise KVWaitTimeout(bad)


class KVCacheCommManager:
    """Bootstrap + registration (reference comm_utils.py:150-222).

    Cross-process (one process per GPU): handles travel through `torch.distributed`
    (`group`, default = the world group the worker already created, reference worker.py:427-432);
    `init_method` ("tcp://ip:port") is only used to create a gloo group when none exists -- it takes
    the place of the reference's MSCCL++ bootstrap address.
    Single process: pass `peer_kv_cache` (+ optional `peer_flags`) to `setup_comm` to wire this rank
    to tensors of another GPU (peer access) or of the same GPU (loopback, tests).
    """

    def __init__(self, rank: int, world_size: int, num_prompt_workers: int,
                 init_method: Optional[str] = None, group=None, device: Optional[torch.device] = None) -> None:
        self.rank = rank
        self.world_size = world_size
        self.num_prompt_workers = num_prompt_workers
        self.corr_worker_rank = pair_rank(rank, world_size, num_prompt_workers)
        self.is_prompt = is_prompt_rank(rank, num_prompt_workers)
        self.init_method = init_method
        self.group = group
        self.device = torch.device(device) if device is not None else None
        self.kvcache_comm: Optional[KVCacheCommunicator] = None
        self._opened: List[int] = []
        self._keepalive: List[torch.Tensor] = []

    # -- helpers ---------------------------------------------------------------------------
    def _export(self, t: torch.Tensor) -> Tuple[bytes, int]:
        handle = (ctypes.c_ubyte * _lib.IPC_HANDLE_BYTES)()
        off, size = ctypes.c_int64(), ctypes.c_int64()
        st = _lib.load().b200_kv_ipc_export(t.data_ptr(), handle, ctypes.byref(off), ctypes.byref(size))
        _lib.check(st, "kv ipc export")
        return bytes(handle), off.value

    def _open(self, handle: bytes, cache: dict) -> int:
        if handle not in cache:  # several tensors can live in one allocation: map it once
            buf = (ctypes.c_ubyte * _lib.IPC_HANDLE_BYTES).from_buffer_copy(handle)
            base = ctypes.c_void_p()
            st = _lib.load().b200_kv_ipc_open(buf, ctypes.byref(base))
            _lib.check(st, "kv ipc open")
            cache[handle] = base.value
            self._opened.append(base.value)
        return cache[handle]

    def _ensure_group(self):
        import torch.distributed as dist
        if self.group is not None:
            return self.group
        if not dist.is_initialized():
            if not self.init_method:
                raise RuntimeError("torch.distributed is not initialised and no init_method was given")
            dist.init_process_group("gloo", init_method=self.init_method, rank=self.rank,
                                    world_size=self.world_size)
        return dist.group.WORLD

    # -- reference surface -------------------------------------------------------------------
    def setup_comm(self, num_layers: int, kv_cache: List[Tuple[torch.Tensor, torch.Tensor]],
                   peer_kv_cache: Optional[List[Tuple[torch.Tensor, torch.Tensor]]] = None,
                   peer_flags: Optional[torch.Tensor] = None, max_ctas: int = 0, defer_puts: bool = True,
                   use_comm_stream: bool = False, layers_per_launch: int = 1) -> None:
        device = kv_cache[0][0].device if self.device is None else self.device
        self.device = device
        k0 = kv_cache[0][0]
        block_size = (k0.numel() * k0.element_size()) // k0.size(0)  # bytes (comm_utils.py:199-201)
        local_flags = torch.zeros(MAX_SLOT_IDS, dtype=torch.int64, device=device)
        bufs = [kv_cache[l][ht] for l in range(num_layers) for ht in HEAD_TYPES]
        src_table = torch.tensor([t.data_ptr() for t in bufs], dtype=torch.int64, device=device)
        dst_table, remote_flags = None, None

        if peer_kv_cache is not None:
            # single-process wiring (another GPU of this process, or loopback on the same GPU)
            peer_dev = peer_kv_cache[0][0].device
            if peer_dev != device:
                st = _lib.load().b200_kv_enable_peer_access(device.index, peer_dev.index)
                _lib.check(st, "kv enable peer access")
            pbufs = [peer_kv_cache[l][ht] for l in range(num_layers) for ht in HEAD_TYPES]
            dst_table = torch.tensor([t.data_ptr() for t in pbufs], dtype=torch.int64, device=device)
            flags = peer_flags if peer_flags is not None else local_flags
            remote_flags = flags.data_ptr()
            self._keepalive += [flags] + pbufs
        else:
            import torch.distributed as dist
            group = self._ensure_group()
            mine = {"rank": self.rank, "bufs": [self._export(t) for t in bufs],
                    "flags": self._export(local_flags)}
            peer = exchange_peer_info(group, self.rank, self.world_size, self.num_prompt_workers, mine)
            if peer is not None:
                opened: dict = {}
                with torch.cuda.device(device):
                    addrs = [self._open(h, opened) + off for h, off in peer["bufs"]]
                    fh, foff = peer["flags"]
                    remote_flags = self._open(fh, opened) + foff
                dst_table = torch.tensor(addrs, dtype=torch.int64, device=device)
            dist.barrier(group=group)

        self.kvcache_comm = KVCacheCommunicator(block_size, num_layers, device, src_table, dst_table, remote_flags,
                                                local_flags, self.rank, self.corr_worker_rank, max_ctas=max_ctas,
                                                defer_puts=defer_puts,
                                                comm_stream=torch.cuda.Stream(device) if use_comm_stream else None,
                                                layers_per_launch=layers_per_launch)
        self.put = self.kvcache_comm.put  # the hot call of the prefill loop: one Python frame, not two

    def flush(self) -> None:
        """Launch any deferred puts now (optional: call after a layer's put loop for the earliest start)."""
        self.kvcache_comm.flush()

    def join_current_stream(self) -> None:
        """Side stream only: the caller's stream waits for every put / signal issued so far."""
        self.kvcache_comm.join_current_stream()

    def destroy_comm(self) -> None:
        if self.kvcache_comm is not None:
            self.kvcache_comm.flush()
        if self.device is not None and self.device.type == "cuda":
            torch.cuda.synchronize(self.device)
        for base in self._opened:
            _lib.load().b200_kv_ipc_close(base)
        self._opened = []
        self._keepalive = []
        self.kvcache_comm = None
        self.__dict__.pop("put", None)

    def wait(self, sem_id: int) -> None:
        self.kvcache_comm.wait(sem_id)

    def check_timeouts(self) -> None:
        self.kvcache_comm.check_timeouts()

    def reset_semaphore(self, sem_id: int) -> None:
        self.kvcache_comm.reset_semaphore(sem_id)

    def signal_and_flush(self, sem_id: int) -> None:
        self.kvcache_comm.signal_and_flush(sem_id)

    def put(self, sem_id: int, layer_id: int, block_start: int, num_blocks: int) -> None:
        self.kvcache_comm.put(sem_id, layer_id, block_start, num_blocks)  # setup_comm rebinds self.put to skip this frame

    def put_runs(self, sem_id: int, layer_id: int, runs: Sequence[Tuple[int, int]]) -> None:
        self.kvcache_comm.put_runs(sem_id, layer_id, runs)

    def put_layers(self, sem_id: int, first_layer: int, num_layers: int, runs: Sequence[Tuple[int, int]]) -> None:
        self.kvcache_comm.put_layers(sem_id, first_layer, num_layers, runs)
