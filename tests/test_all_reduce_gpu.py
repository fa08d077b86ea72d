"""GPU parity of the one-shot NVLink all-reduce against oracle.all_reduce_sum: bit-exact (fp32
accumulation in rank order, one rounding), identical on every rank.  All ranks live in this process (one per GPU,
peer access); the multi-GPU cases are skipped on a one-GPU box, where the world-size-1 case still exercises the
kernel, the flag protocol and the C ABI."""
import pytest
import torch

from oracle import paged_kv_oracle as O
from _util import bits

pytestmark = pytest.mark.gpu


def _comms(world):
    from vllm_splitwise_b200.custom_all_reduce import CustomAllreduce
    comms = [CustomAllreduce(r, world, f"cuda:{r}", max_size=16 << 20) for r in range(world)]
    CustomAllreduce.connect_local(comms)
    return comms


def _run(world, numel, dtype, calls=3, two_stage_min_bytes=None):
    comms = _comms(world)
    for c in comms:
        if two_stage_min_bytes is not None:
            c.two_stage_min_bytes = two_stage_min_bytes
    for call in range(calls):
        g = torch.Generator().manual_seed(100 * call + world)
        xs = [(torch.rand(numel, generator=g) * 4 - 2).to(dtype) for _ in range(world)]
        outs = []
        for r in range(world):  # launches are asynchronous: the per-GPU kernels run concurrently and meet at the flags
            with torch.cuda.device(r):
                outs.append(comms[r].all_reduce(xs[r].to(f"cuda:{r}")))
        ref = O.all_reduce_sum(xs)
        for r in range(world):
            torch.cuda.synchronize(r)
            assert not comms[r].timed_out()
            assert torch.equal(bits(outs[r].cpu()), bits(ref)), f"rank {r} call {call}"


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16, torch.float32])
@pytest.mark.parametrize("numel", [8, 4096, 64 * 4096 + 8])
def test_world1(dtype, numel):
    _run(1, numel, dtype)


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16, torch.float32])
def test_multi_gpu(world, dtype):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    for numel in (8, 64 * 4096, 256 * 8192):   # C2 and C3 decode activations [batch, hidden]
        _run(world, numel, dtype)


@pytest.mark.parametrize("world", [4, 8])
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16, torch.float32])
def test_two_stage_multi_gpu(world, dtype):
    """reduce-scatter + all-gather path, forced for every size (ragged ranges included): same bits as the oracle"""
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    for numel in (8, 8 * 37, 64 * 4096 + 24, 2048 * 1280):
        _run(world, numel, dtype, calls=2, two_stage_min_bytes=0)


def test_registered_input_skips_the_staging_copy():
    c = _comms(1)[0]
    x = c.registered_input((4, 64), torch.float16)
    x.copy_(torch.arange(256, dtype=torch.float16).view(4, 64))
    out = c.all_reduce(x)
    torch.cuda.synchronize()
    assert torch.equal(out, x) and x.data_ptr() == c.buffer.data_ptr()


def test_argument_errors():
    c = _comms(1)[0]
    with pytest.raises(RuntimeError):
        c.all_reduce(torch.zeros(3, dtype=torch.float16, device="cuda:0"))      # 6 bytes: not a multiple of 16
    with pytest.raises(RuntimeError):
        c.all_reduce(torch.zeros(8, dtype=torch.float16))                       # CPU tensor
    with pytest.raises(RuntimeError, match="Unsupported data type"):
        c.all_reduce(torch.zeros(16, dtype=torch.int32, device="cuda:0"))
