"""world_size-2 `gloo` tests (CPU) of the N > 1 host logic: the KV-head shard each rank takes, the
bootstrap exchange of the KV transfer, and bench.py's max-over-ranks reduction.  The device kernels are
covered by the -m gpu tests; here every rank runs the ORACLE on its shard (checker only) to show that
the partition composes to the unsharded result with no exchange inside the path."""
import os
import sys
import tempfile

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, initfile, fn_name, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    torch.set_num_threads(1)
    dist.init_process_group("gloo", init_method=f"file://{initfile}", rank=rank, world_size=world)
    try:
        globals()[fn_name](rank, world)
        q.put((rank, "ok"))
    except Exception as e:  # noqa: BLE001
        import traceback
        q.put((rank, "".join(traceback.format_exception(e))))
    finally:
        dist.destroy_process_group()


def _run(fn_name, world=2):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    with tempfile.TemporaryDirectory() as d:
        initfile = os.path.join(d, "init")
        procs = [ctx.Process(target=_worker, args=(r, world, initfile, fn_name, q)) for r in range(world)]
        for p in procs:
            p.start()
        results = [q.get(timeout=240) for _ in procs]
        for p in procs:
            p.join(timeout=60)
    for rank, res in results:
        assert res == "ok", f"rank {rank}: {res}"


# ---- rank bodies ---------------------------------------------------------------------------
def _body_head_shards_compose(rank, world):
    from oracle import paged_kv_oracle as O
    from vllm_splitwise_b200.sharding import shard_heads
    torch.manual_seed(0)  # same inputs on every rank (block tables etc. are replicated, model_runner.py:525-547)
    S, Hq, Hkv, D, bs, NB = 3, 8, 2, 64, 16, 12
    scale = D ** -0.5
    x = 4
    q = torch.empty(S, Hq, D).uniform_(-scale, scale)
    kc = torch.empty(NB, Hkv, D // x, bs, x).uniform_(-scale, scale)
    vc = torch.empty(NB, Hkv, D, bs).uniform_(-scale, scale)
    bt = torch.stack([torch.randperm(NB)[:4] for _ in range(S)]).to(torch.int32)
    cl = torch.tensor([1, 33, 64], dtype=torch.int32)
    sh = shard_heads(Hq, Hkv, world, rank)
    q_r = q[:, sh.q_start:sh.q_start + sh.num_q].contiguous()
    kc_r = kc[:, sh.kv_start:sh.kv_start + sh.num_kv].contiguous()
    vc_r = vc[:, sh.kv_start:sh.kv_start + sh.num_kv].contiguous()
    out_r = torch.empty_like(q_r)
    O.single_query_cached_kv_attention(out_r, q_r, sh.num_q // sh.num_kv, kc_r, vc_r, bt, cl, scale, None)
    parts = [torch.empty_like(out_r) for _ in range(world)]
    dist.all_gather(parts, out_r)  # test-side gather only: the path itself has no collective
    full = torch.empty_like(q)
    O.single_query_cached_kv_attention(full, q, Hq // Hkv, kc, vc, bt, cl, scale, None)
    assert torch.equal(torch.cat(parts, dim=1), full)

    # reshape_and_cache shards the same way: each rank writes its heads, slots replicated
    key = torch.empty(S, Hkv, D).uniform_(-1, 1)
    val = torch.empty(S, Hkv, D).uniform_(-1, 1)
    slots = torch.tensor([5, -1, 40], dtype=torch.int64)
    kc_full, vc_full = kc.clone(), vc.clone()
    O.reshape_and_cache(key, val, kc_full, vc_full, slots, "auto")
    O.reshape_and_cache(key[:, sh.kv_start:sh.kv_start + sh.num_kv].contiguous(),
                        val[:, sh.kv_start:sh.kv_start + sh.num_kv].contiguous(), kc_r, vc_r, slots, "auto")
    assert torch.equal(kc_r, kc_full[:, sh.kv_start:sh.kv_start + sh.num_kv])
    assert torch.equal(vc_r, vc_full[:, sh.kv_start:sh.kv_start + sh.num_kv])


def _body_kv_bootstrap_exchange(rank, world):
    from vllm_splitwise_b200.comm_utils import exchange_peer_info, is_prompt_rank, pair_rank
    npw = world // 2
    mine = {"rank": rank, "bufs": [(bytes([rank]) * 64, 4096 * i) for i in range(4)], "flags": (bytes([100 + rank]) * 64, 0)}
    peer = exchange_peer_info(dist.group.WORLD, rank, world, npw, mine)
    if is_prompt_rank(rank, npw):
        want = pair_rank(rank, world, npw)
        assert peer["rank"] == want and peer["bufs"][2] == (bytes([want]) * 64, 8192)
        assert peer["flags"][0] == bytes([100 + want]) * 64
    else:
        assert peer is None


def _body_bench_max_over_ranks(rank, world):
    import bench
    got = bench.max_over_ranks([1.0 + rank, 10.0 - rank, 5.0], torch.device("cpu"), world)
    assert got == [float(world), 10.0, 5.0]


# ---- tests ---------------------------------------------------------------------------------
def test_head_shards_compose_world2():
    _run("_body_head_shards_compose", 2)


def test_kv_bootstrap_exchange_world2():
    _run("_body_kv_bootstrap_exchange", 2)


def test_kv_bootstrap_exchange_world4():
    _run("_body_kv_bootstrap_exchange", 4)


def test_bench_max_over_ranks_world2():
    _run("_body_bench_max_over_ranks", 2)
