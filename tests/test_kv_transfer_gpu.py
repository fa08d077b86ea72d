"""GPU tests of the Splitwise prefill -> decode KV-block transfer (KVCacheCommManager surface,
reference vllm/worker/comm_utils.py:150-222; functional model tests/distributed/test_kvcache_comm.py and
vllm/worker/worker.py:350-386).  Oracle: "the same bytes at the same block offsets, nothing else touched"
(oracle.kv_put).  One-GPU cases wire the manager to a second set of cache tensors on the same GPU
(loopback); the two-GPU case uses peer access and is skipped on a one-GPU box.  Fragmented and
multi-sequence transfers are NOT pinned by any reference test (parity unpinned).

No test here lets a wait kernel spin on a flag that a later launch on the same GPU would set: waits are
only enqueued after the matching signal on the same stream (or with a tiny spin bound for the timeout test).
"""
import random

import pytest
import torch

from _util import bits

pytestmark = pytest.mark.gpu


def _caches(num_layers, num_blocks, hkv, d, bs, dtype, device, seed):
    g = torch.Generator(device=device).manual_seed(seed)
    x = 16 // torch.tensor([], dtype=dtype).element_size()
    out = []
    for _ in range(num_layers):
        k = torch.randint(-30000, 30000, (num_blocks, hkv, d // x, bs, x), dtype=torch.int16, device=device, generator=g)
        v = torch.randint(-30000, 30000, (num_blocks, hkv, d, bs), dtype=torch.int16, device=device, generator=g)
        out.append((k.view(dtype) if dtype != torch.int16 else k, v.view(dtype) if dtype != torch.int16 else v))
    return out


def _manager(src, dst, num_layers, max_ctas=0):
    from vllm_splitwise_b200.comm_utils import KVCacheCommManager
    m = KVCacheCommManager(rank=0, world_size=2, num_prompt_workers=1)
    m.setup_comm(num_layers, src, peer_kv_cache=dst, max_ctas=max_ctas)
    return m


def _check(src, dst, dst_before, blocks_by_layer):
    """dst == src on the transferred blocks, dst unchanged elsewhere (both K and V, every layer)."""
    for l, blocks in blocks_by_layer.items():
        moved = torch.zeros(src[l][0].shape[0], dtype=torch.bool, device=src[l][0].device)
        moved[torch.tensor(sorted(blocks), device=moved.device)] = True
        for kv in (0, 1):
            s, d, b = bits(src[l][kv]), bits(dst[l][kv]), bits(dst_before[l][kv])
            assert torch.equal(d[moved], s[moved])
            assert torch.equal(d[~moved], b[~moved])


def test_put_signal_wait_like_reference_functional_test():
    """worker.py:350-386: prompt side fills its cache with a constant, puts ALL blocks of every layer under
    one semaphore, signals; token side waits and sees the constant everywhere."""
    dev = torch.device("cuda:0")
    L, NB = 4, 64
    src = [(torch.full((NB, 10, 16, 16, 8), 1.5, dtype=torch.half, device=dev),
            torch.full((NB, 10, 128, 16), 1.5, dtype=torch.half, device=dev)) for _ in range(L)]
    dst = [(torch.zeros_like(k), torch.zeros_like(v)) for k, v in src]
    m = _manager(src, dst, L)
    assert m.kvcache_comm.block_size == 10 * 128 * 16 * 2
    for layer in range(L):
        m.put(0, layer, 0, NB)
    m.signal_and_flush(0)
    m.wait(0)
    torch.cuda.synchronize()
    assert not m.kvcache_comm.timed_out()
    for k, v in dst:
        assert bool((k == 1.5).all()) and bool((v == 1.5).all())
    m.destroy_comm()


@pytest.mark.parametrize("dtype", [torch.half, torch.uint8])
@pytest.mark.parametrize("pattern", ["contiguous", "fragmented", "mixed"])
@pytest.mark.parametrize("max_ctas", [0, 8])
def test_block_lists_bit_exact(dtype, pattern, max_ctas):
    from oracle import paged_kv_oracle as O
    from vllm_splitwise_b200.utils import coalesce_blocks
    random.seed(0)
    dev = torch.device("cuda:0")
    L, NB, hkv, d, bs = 6, 300, 10, 128, 16
    if dtype == torch.uint8:
        src = [(torch.randint(0, 256, (NB, hkv, d // 16, bs, 16), dtype=torch.uint8, device=dev),
                torch.randint(0, 256, (NB, hkv, d, bs), dtype=torch.uint8, device=dev)) for _ in range(L)]
    else:
        src = _caches(L, NB, hkv, d, bs, dtype, dev, 0)
    dst = [(torch.randint_like(k.view(torch.uint8), 256).view(k.dtype).view(k.shape),
            torch.randint_like(v.view(torch.uint8), 256).view(v.dtype).view(v.shape)) for k, v in src]
    before = [(k.clone(), v.clone()) for k, v in dst]
    if pattern == "contiguous":
        blocks = list(range(37, 37 + 128))
    elif pattern == "fragmented":
        blocks = random.sample(range(0, NB, 2), 128)   # no two ids adjacent: 128 runs of one block
    else:
        blocks = list(range(5, 40)) + random.sample(range(50, NB), 60)
    runs = coalesce_blocks(blocks)
    assert [tuple(r) for r in runs] == O.coalesce_blocks(blocks)
    m = _manager(src, dst, L, max_ctas)
    # layer 0 by the reference's per-run call, layers 1-2 per layer, layers 3-5 in one launch
    for start, n in runs:
        m.put(3, 0, start, n)
    m.put_runs(3, 1, runs)
    m.put_runs(3, 2, runs)
    m.put_layers(3, 3, 3, runs)
    m.signal_and_flush(3)
    m.wait(3)
    torch.cuda.synchronize()
    assert not m.kvcache_comm.timed_out()
    _check(src, dst, before, {l: blocks for l in range(L)})
    # the oracle on the same inputs gives the same destination image
    want = [(k.clone(), v.clone()) for k, v in before]
    O.kv_put([t for kv in src for t in kv], [t for kv in want for t in kv], runs)
    for (k, v), (wk, wv) in zip(dst, want):
        assert torch.equal(bits(k), bits(wk)) and torch.equal(bits(v), bits(wv))
    m.destroy_comm()


@pytest.mark.parametrize("defer", [True, False])
@pytest.mark.parametrize("side_stream", [False, True])
def test_reference_call_pattern_many_semaphores(defer, side_stream):
    """The backends' loop (flash_attn.py:81-86 + worker.py:327-331): per layer, per semaphore, per coalesced
    run one put(); signals after the last layer.  Deferred puts leave as one launch per layer; eager puts as
    one launch each; with a side stream the transfers are ordered after the caller's stream by an event."""
    from vllm_splitwise_b200.comm_utils import KVCacheCommManager
    from vllm_splitwise_b200.utils import coalesce_blocks
    random.seed(11)
    dev = torch.device("cuda:0")
    L, NB, hkv, d, bs = 5, 400, 10, 128, 16
    src = _caches(L, NB, hkv, d, bs, torch.half, dev, 4)
    dst = [(torch.zeros_like(k), torch.zeros_like(v)) for k, v in src]
    before = [(k.clone(), v.clone()) for k, v in dst]
    ids = random.sample(range(300), 240)
    per_sem = {s: ids[s * 60:(s + 1) * 60] for s in range(4)}     # 4 prompts in the batch, mixed run lengths
    per_sem[1] = list(range(300, 360))                            # one of them a single contiguous run
    m = KVCacheCommManager(rank=0, world_size=2, num_prompt_workers=1)
    m.setup_comm(L, src, peer_kv_cache=dst, defer_puts=defer, use_comm_stream=side_stream)
    lib = m.kvcache_comm._lib
    n0 = lib.b200_kv_launch_count()
    total_runs = 0
    for layer in range(L):
        for sem, blocks in per_sem.items():
            for start, n in coalesce_blocks(blocks):
                m.put(sem, layer, start, n)
                total_runs += 1
    for sem in per_sem:
        m.signal_and_flush(sem)
    put_launches = lib.b200_kv_launch_count() - n0 - len(per_sem)
    if side_stream:   # the consumer's stream must be ordered after the side stream's signals
        torch.cuda.current_stream().wait_stream(m.kvcache_comm.comm_stream)
    for sem in per_sem:
        m.wait(sem)
    torch.cuda.synchronize()
    assert not m.kvcache_comm.timed_out()
    assert put_launches == (L if defer else total_runs), (put_launches, total_runs)
    moved = sorted(b for bl in per_sem.values() for b in bl)
    _check(src, dst, before, {l: moved for l in range(L)})
    m.destroy_comm()


def test_deferred_puts_flush_points():
    """A deferred put is launched by: a put for another layer, flush(), signal, put_runs/put_layers, destroy."""
    dev = torch.device("cuda:0")
    src = _caches(3, 32, 2, 64, 16, torch.half, dev, 6)
    dst = [(torch.zeros_like(k), torch.zeros_like(v)) for k, v in src]
    m = _manager(src, dst, 3)
    c = m.kvcache_comm
    m.put(0, 0, 0, 4)
    torch.cuda.synchronize()
    assert not bool(dst[0][0][:4].any())            # still on the host
    m.put(0, 1, 4, 4)                               # layer change -> layer 0 leaves
    torch.cuda.synchronize()
    assert torch.equal(bits(dst[0][0][:4]), bits(src[0][0][:4])) and not bool(dst[1][0].any())
    m.flush()
    torch.cuda.synchronize()
    assert torch.equal(bits(dst[1][1][4:8]), bits(src[1][1][4:8]))
    m.put(0, 2, 8, 2)
    m.put(0, 2, 20, 3)
    assert c._pending_runs == [(8, 2), (20, 3)]
    m.signal_and_flush(0)
    m.wait(0)
    torch.cuda.synchronize()
    assert c._pending_runs == [] and torch.equal(bits(dst[2][0][20:23]), bits(src[2][0][20:23]))
    assert not bool(dst[2][0][10:20].any())
    with pytest.raises(ValueError):
        m.put(0, 3, 0, 1)                           # layer out of range
    m.put(0, 0, 30, 0)                              # empty run: nothing queued
    assert c._pending_runs == []
    m.destroy_comm()


def test_many_runs_exceeding_one_launch_table():
    """More runs than one kernel-parameter table holds (the library splits the launch)."""
    dev = torch.device("cuda:0")
    L, NB = 2, 2048
    src = _caches(L, NB, 2, 64, 16, torch.half, dev, 1)
    dst = [(torch.zeros_like(k), torch.zeros_like(v)) for k, v in src]
    before = [(k.clone(), v.clone()) for k, v in dst]
    blocks = list(range(0, NB, 2))  # 1024 single-block runs
    m = _manager(src, dst, L)
    m.put_layers(0, 0, L, [(b, 1) for b in blocks])
    m.signal_and_flush(0)
    m.wait(0)
    torch.cuda.synchronize()
    _check(src, dst, before, {0: blocks, 1: blocks})
    m.destroy_comm()


def test_semaphores_are_counters_and_wait_times_out():
    dev = torch.device("cuda:0")
    src = _caches(1, 8, 1, 64, 16, torch.half, dev, 2)
    dst = [(torch.zeros_like(k), torch.zeros_like(v)) for k, v in src]
    m = _manager(src, dst, 1)
    for _ in range(3):  # three rounds on one semaphore, one on another
        m.put(7, 0, 0, 8)
        m.signal_and_flush(7)
        m.wait(7)
    m.put(9, 0, 0, 1)
    m.signal_and_flush(9)
    m.wait(9)
    torch.cuda.synchronize()
    assert not m.kvcache_comm.timed_out()
    flags = m.kvcache_comm.local_flags.cpu()
    assert int(flags[7]) == 3 and int(flags[9]) == 1 and int(flags.sum()) == 4
    # a wait nobody signals: bounded spin, reports the timeout instead of hanging
    m.kvcache_comm.wait_max_spins = 64
    m.wait(11)
    torch.cuda.synchronize()
    assert m.kvcache_comm.timed_out()
    with pytest.raises(ValueError):
        m.put(256, 0, 0, 1)
    # The timeout is not swallowed: the step-boundary check raises, names the semaphore and resynchronises it
    # (without that, `expected` would stay one ahead and every later wait on semaphore 11 would spin out).
    from vllm_splitwise_b200.comm_utils import KVWaitTimeout
    with pytest.raises(KVWaitTimeout) as ei:
        m.check_timeouts()
    assert ei.value.sem_ids == [11]
    assert not m.kvcache_comm.timed_out()
    m.check_timeouts()  # nothing pending any more
    # signal-after-timeout: the late transfer of that sequence, then the next wait, complete normally
    m.kvcache_comm.wait_max_spins = 100000000
    m.put(11, 0, 2, 3)
    m.signal_and_flush(11)
    m.wait(11)
    torch.cuda.synchronize()
    assert not m.kvcache_comm.timed_out()
    assert int(m.kvcache_comm.local_flags[11]) == 1 and m.kvcache_comm.expected[11] == 1
    # an aborted sequence: the signal arrives but the token side never waits for it -> reset_semaphore realigns
    m.signal_and_flush(11)
    torch.cuda.synchronize()
    m.reset_semaphore(11)
    assert m.kvcache_comm.expected[11] == 2
    m.signal_and_flush(11)
    m.wait(11)
    torch.cuda.synchronize()
    assert not m.kvcache_comm.timed_out()
    m.destroy_comm()


def test_two_gpu_peer_scatter():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    from vllm_splitwise_b200.comm_utils import KVCacheCommManager
    from vllm_splitwise_b200.utils import MAX_SLOT_IDS, coalesce_blocks
    random.seed(5)
    d0, d1 = torch.device("cuda:0"), torch.device("cuda:1")
    L, NB = 8, 256
    src = _caches(L, NB, 10, 128, 16, torch.half, d0, 3)
    dst = [(torch.zeros_like(k, device=d1), torch.zeros_like(v, device=d1)) for k, v in src]
    before = [(k.clone(), v.clone()) for k, v in dst]
    peer_flags = torch.zeros(MAX_SLOT_IDS, dtype=torch.int64, device=d1)
    m = KVCacheCommManager(rank=0, world_size=2, num_prompt_workers=1)
    m.setup_comm(L, src, peer_kv_cache=dst, peer_flags=peer_flags)
    blocks = random.sample(range(NB), 128)
    m.put_layers(1, 0, L, coalesce_blocks(blocks))
    m.signal_and_flush(1)
    torch.cuda.synchronize(d0)
    torch.cuda.synchronize(d1)
    assert int(peer_flags[1]) == 1
    src_on_1 = [(k.to(d1), v.to(d1)) for k, v in src]
    _check(src_on_1, dst, before, {l: blocks for l in range(L)})
    m.destroy_comm()


def test_two_process_ipc_transfer_bit_exact():
    """One process per GPU (the deployment shape): handles cross processes by CUDA IPC, the prompt rank's
    kernel stores into the token rank's cache, release/acquire flags order it.  Runs the C5 harness at a
    small size; it verifies the received blocks bit for bit and exits non-zero on any mismatch."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import json
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", "29631", os.path.join(root, "tools", "kv_transfer_bench.py"),
           "--prompts", "3", "--layers", "4", "--prompt-tokens", "512", "--iters", "2", "--warmup", "1"]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-4000:]
    lines = [json.loads(l) for l in res.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 8 and all(l["bit_exact"] for l in lines)


@pytest.fixture
def put_config():
    """Set / restore the process-wide put configuration (b200_kv_put_configure)."""
    import ctypes
    from vllm_splitwise_b200 import _lib
    lib = _lib.load()
    v, c = ctypes.c_int32(), ctypes.c_int32()
    assert lib.b200_kv_put_config(ctypes.byref(v), ctypes.byref(c)) == 0
    yield lib
    assert lib.b200_kv_put_configure(v.value, c.value) == 0


@pytest.mark.parametrize("variant", [0, 1])
@pytest.mark.parametrize("ctas", [1, 5, 32, 0])
@pytest.mark.parametrize("layers_per_launch", [1, 3])
def test_put_variants_cta_budgets_and_layer_batches(put_config, variant, ctas, layers_per_launch):
    """Both put kernels (SIMT warp chunks / cp.async.bulk ring), any CTA budget, layers merged into one launch:
    the same bytes arrive.  Block size 40 960 B (two 20 KiB tiles, ten 4 KiB chunks), mixed run lengths, and a
    second prompt whose run list is longer than one parameter table (device-table kernels)."""
    from vllm_splitwise_b200.comm_utils import KVCacheCommManager
    from vllm_splitwise_b200.utils import coalesce_blocks
    assert put_config.b200_kv_put_configure(variant, 0) == 0
    assert put_config.b200_kv_put_configure(7, 0) == -4   # unknown variant: B200_ERR_ARG, configuration unchanged
    random.seed(5)
    dev = torch.device("cuda:0")
    L, NB, hkv, d, bs = 7, 1400, 10, 128, 16
    src = _caches(L, NB, hkv, d, bs, torch.half, dev, 9)
    dst = [(torch.zeros_like(k), torch.zeros_like(v)) for k, v in src]
    before = [(k.clone(), v.clone()) for k, v in dst]
    a = list(range(3, 60)) + random.sample(range(70, 300), 90)          # mixed runs
    b = list(range(301, NB, 2))                                         # 550 single-block runs (> 440)
    m = KVCacheCommManager(rank=0, world_size=2, num_prompt_workers=1)
    m.setup_comm(L, src, peer_kv_cache=dst, max_ctas=ctas, layers_per_launch=layers_per_launch)
    lib = m.kvcache_comm._lib
    n0 = lib.b200_kv_launch_count()
    ra, rb = tuple(map(tuple, coalesce_blocks(a))), tuple(map(tuple, coalesce_blocks(b)))
    for layer in range(L):
        m.put_runs(0, layer, ra)
    m.signal_and_flush(0)
    launches_a = lib.b200_kv_launch_count() - n0 - 1
    assert launches_a == -(-L // layers_per_launch)
    for layer in range(L - 1):                  # layer L-1 of prompt b is NOT sent
        for s, n in rb:
            m.put(1, layer, s, n)
    m.signal_and_flush(1)
    m.wait(0)
    m.wait(1)
    torch.cuda.synchronize()
    assert not m.kvcache_comm.timed_out()
    want = {l: sorted(a + b) for l in range(L - 1)}
    want[L - 1] = sorted(a)
    _check(src, dst, before, want)
    m.destroy_comm()
