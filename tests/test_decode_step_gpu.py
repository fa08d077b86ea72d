"""The composed decode step (b200_decode_step / decode_step.DecodeStepRunner): device-side metadata -> RoPE fused
into the cache write -> paged attention, chained with programmatic dependent launch and captured as a CUDA graph
per batch size (reference: vllm/worker/model_runner.py:274-399, 826-873; models/llama.py:156-158;
attention/ops/paged_attn.py:21-111).  Checked, step by step over a running batch that crosses block boundaries,
against the three oracles chained: prepare_decode -> rotary_embedding + reshape_and_cache -> attention.
Metadata, rotated q / k and the caches are BIT-EXACT; the attention output meets the attention tolerance."""
import random

import pytest
import torch

from oracle import paged_kv_oracle as O
from _util import bits

pytestmark = pytest.mark.gpu


def _setup(dtype, kv, Hq, Hkv, D, bs, B, pad_to, steps, seed, ctx0):
    from vllm_splitwise_b200.decode_metadata import DeviceDecodeState
    from vllm_splitwise_b200.decode_step import DecodeStepRunner
    from vllm_splitwise_b200.utils import create_kv_caches_with_random
    random.seed(seed)
    torch.manual_seed(seed)
    dev = torch.device("cuda")
    max_blocks = (max(ctx0) + steps) // bs + 2
    NB = B * max_blocks + 3
    layers = 2
    kcs, vcs = create_kv_caches_with_random(NB, bs, layers, Hkv, D, kv, dtype, seed, dev)
    free = list(range(NB))
    random.shuffle(free)
    tables = [[free.pop() for _ in range((c + bs - 1) // bs)] for c in ctx0]   # blocks owned before the first step
    st = DeviceDecodeState(B + 3, max_blocks, bs, dev, max_batch=pad_to)
    rows = random.sample(range(B + 3), B)
    for r, t, c in zip(rows, tables, ctx0):
        st.add_sequence(r, t, c)
    st.set_batch(rows, pad_to=pad_to)
    cos_sin = O.compute_cos_sin_cache(D, max(ctx0) + steps + 8, 10000.0, dtype).to(dev)
    runner = DecodeStepRunner(st, Hq, Hkv, D, D ** -0.5, dtype, kv, cos_sin, True, None)
    return dict(st=st, runner=runner, kcs=kcs, vcs=vcs, tables=tables, lens=list(ctx0), rows=rows, free=free,
                cos_sin=cos_sin, dev=dev, layers=layers)


def _oracle_step(env, qkv_host, Hq, Hkv, D, bs, kv, dtype, pad_to, max_blocks, ref_kcs, ref_vcs):
    """One step of the chained oracles on CPU copies; returns (positions, slots, ctx, tables, q_rot, k_rot, out)."""
    B = len(env["lens"])
    lens = [n + 1 for n in env["lens"]]                       # length including the token being fed
    pos, slots, ctx, bt = O.prepare_decode(lens, env["tables"], bs, None, pad_to=pad_to, out_blocks=max_blocks)
    outs, qs, ks = [], [], []
    for layer in range(env["layers"]):
        row = qkv_host[layer].clone()
        q, k, v = row.split([Hq * D, Hkv * D, Hkv * D], dim=-1)
        O.rotary_embedding_reshape_and_cache(pos.flatten(), q, k, v, D, env["cos_sin"].cpu(), True, ref_kcs[layer],
                                             ref_vcs[layer], slots.flatten(), kv)
        kc, vc = ref_kcs[layer], ref_vcs[layer]
        if kv == "fp8_e5m2":
            kc, vc = O.e5m2_to_tensor(kc, dtype), O.e5m2_to_tensor(vc, dtype)
        out = torch.zeros(pad_to, Hq, D, dtype=dtype)
        O.single_query_cached_kv_attention(out, q.reshape(pad_to, Hq, D), Hq // Hkv, kc, vc, bt, ctx, D ** -0.5, None)
        outs.append(out); qs.append(q.clone()); ks.append(k.clone())
    return pos, slots, ctx, bt, qs, ks, outs


@pytest.mark.parametrize("dtype,kv,heads", [(torch.half, "auto", (8, 8)), (torch.bfloat16, "auto", (16, 2)),
                                            (torch.bfloat16, "fp8_e5m2", (32, 8)), (torch.float, "auto", (4, 4))])
@pytest.mark.parametrize("captured", [True, False])
def test_decode_step_matches_chained_oracles(dtype, kv, heads, captured):
    Hq, Hkv = heads
    D, bs, B, pad_to, steps = 128, 16, 5, 8, 20
    ctx0 = [15, 31, 100, 1, 257]          # several cross a block boundary within the first steps
    env = _setup(dtype, kv, Hq, Hkv, D, bs, B, pad_to, steps, 3, ctx0)
    st, runner = env["st"], env["runner"]
    kv_caches = list(zip(env["kcs"], env["vcs"]))
    ref_kcs = [t.cpu().clone() for t in env["kcs"]]
    ref_vcs = [t.cpu().clone() for t in env["vcs"]]
    width = (Hq + 2 * Hkv) * D
    max_ctx = max(ctx0) + steps
    if captured:
        cap = runner.capture(pad_to, kv_caches, max_context_len=max_ctx)
        assert torch.equal(st.seq_lens.cpu()[env["rows"]], torch.tensor(ctx0, dtype=torch.int32))  # warm-up left no trace
        for a, b in zip(env["kcs"] + env["vcs"], ref_kcs + ref_vcs):
            assert torch.equal(bits(a.cpu()), bits(b))
    else:
        qkv_dev = torch.zeros(env["layers"], pad_to, width, dtype=dtype, device=env["dev"])
        out_dev = torch.zeros(env["layers"], pad_to, Hq, D, dtype=dtype, device=env["dev"])
    atol = 1e-2 if kv == "fp8_e5m2" else 1e-3
    for step in range(steps):
        # the block manager hands out a block to every sequence whose new token opens one (block_manager.py:268-309)
        appends = []
        for i, n in enumerate(env["lens"]):
            if n % bs == 0 and n // bs >= len(env["tables"][i]):
                blk = env["free"].pop()
                env["tables"][i].append(blk)
                appends.append((env["rows"][i], n // bs, blk))
        st.append_blocks(appends)
        # inputs by the reference's recipe (uniform(-D^-0.5, D^-0.5), tests/kernels/test_attention.py:137-145): the
        # atol 1e-3 of the attention check is defined on it (one bf16 ulp at |x| ~ 1 is already 3.9e-3)
        qkv_host = torch.empty(env["layers"], pad_to, width, dtype=dtype).uniform_(-D ** -0.5, D ** -0.5)
        if captured:
            cap.qkv_host.copy_(qkv_host)
            cap.replay()
            cap.synchronize()
            out_gpu, qkv_gpu = cap.out_host.clone(), cap.qkv.cpu()
        else:
            qkv_dev.copy_(qkv_host)
            for layer, (kc, vc) in enumerate(kv_caches):
                runner.layer_step(qkv_dev[layer], out_dev[layer], kc, vc, pad_to, max_ctx, with_metadata=layer == 0)
            torch.cuda.synchronize()
            out_gpu, qkv_gpu = out_dev.cpu(), qkv_dev.cpu()
        pos, slots, ctx, bt, qs, ks, outs = _oracle_step(env, qkv_host, Hq, Hkv, D, bs, kv, dtype, pad_to,
                                                         st.max_blocks, ref_kcs, ref_vcs)
        assert torch.equal(st.positions[:pad_to].cpu(), pos), f"step {step}"
        assert torch.equal(st.slot_mapping[:pad_to].cpu(), slots), f"step {step}"
        assert torch.equal(st.context_lens[:pad_to].cpu(), ctx), f"step {step}"
        assert torch.equal(st.out_block_tables[:pad_to].cpu(), bt), f"step {step}"
        for layer in range(env["layers"]):
            q_gpu, k_gpu, _ = qkv_gpu[layer].split([Hq * D, Hkv * D, Hkv * D], dim=-1)
            if dtype == torch.float:
                assert torch.allclose(q_gpu, qs[layer], rtol=1e-6, atol=1e-7) and torch.allclose(k_gpu, ks[layer], rtol=1e-6, atol=1e-7)
            else:
                assert torch.equal(bits(q_gpu), bits(qs[layer])), f"rotated q, step {step} layer {layer}"
                assert torch.equal(bits(k_gpu), bits(ks[layer])), f"rotated k, step {step} layer {layer}"
                assert torch.equal(bits(env["kcs"][layer].cpu()), bits(ref_kcs[layer])), f"key cache, step {step} layer {layer}"
            assert torch.equal(bits(env["vcs"][layer].cpu()), bits(ref_vcs[layer])), f"value cache, step {step} layer {layer}"
            o, r = out_gpu[layer][:B].float(), outs[layer][:B].float()
            assert torch.allclose(o, r, atol=atol, rtol=1e-3), f"attention, step {step} layer {layer}: {(o - r).abs().max()}"
        env["lens"] = [n + 1 for n in env["lens"]]
    assert st.seq_lens.cpu()[env["rows"]].tolist() == env["lens"]


def test_decode_step_v2_branch_and_no_rope():
    """Long context -> the reference rule picks v2 (paged_attn.py:62-63): the composed step carries the v2
    scratch; cos_sin_cache=None composes plain reshape_and_cache instead of the fused RoPE write."""
    from vllm_splitwise_b200.decode_metadata import DeviceDecodeState
    from vllm_splitwise_b200.decode_step import DecodeStepRunner
    from vllm_splitwise_b200.utils import create_kv_caches_with_random
    torch.manual_seed(0)
    dev = torch.device("cuda")
    Hq, Hkv, D, bs, B = 8, 2, 128, 16, 3
    ctx0 = [9000, 8191, 40]
    max_blocks = 9100 // bs
    NB = B * max_blocks
    kcs, vcs = create_kv_caches_with_random(NB, bs, 1, Hkv, D, "auto", torch.half, 0, dev)
    perm = torch.randperm(NB).tolist()
    tables = [perm[i * max_blocks:(i + 1) * max_blocks] for i in range(B)]
    st = DeviceDecodeState(B, max_blocks, bs, dev)
    for r in range(B):
        st.add_sequence(r, tables[r], ctx0[r])
    st.set_batch(list(range(B)), pad_to=B)
    runner = DecodeStepRunner(st, Hq, Hkv, D, D ** -0.5, torch.half, "auto", None, True, None)
    assert runner.v2_scratch(B, 9100) is not None
    cap = runner.capture(B, [(kcs[0], vcs[0])], max_context_len=9100)
    ref_kc, ref_vc = kcs[0].cpu().clone(), vcs[0].cpu().clone()
    for step in range(3):
        qkv = torch.empty(1, B, (Hq + 2 * Hkv) * D, dtype=torch.half).uniform_(-0.1, 0.1)
        cap.qkv_host.copy_(qkv)
        cap.replay()
        cap.synchronize()
        lens = [c + step + 1 for c in ctx0]
        pos, slots, ctx, bt = O.prepare_decode(lens, tables, bs, None, pad_to=B, out_blocks=max_blocks)
        q, k, v = qkv[0].split([Hq * D, Hkv * D, Hkv * D], dim=-1)
        O.reshape_and_cache(k.reshape(B, Hkv, D), v.reshape(B, Hkv, D), ref_kc, ref_vc, slots.flatten(), "auto")
        out = torch.zeros(B, Hq, D, dtype=torch.half)
        O.single_query_cached_kv_attention(out, q.reshape(B, Hq, D), Hq // Hkv, ref_kc, ref_vc, bt, ctx, D ** -0.5, None)
        assert torch.equal(bits(cap.qkv.cpu()[0]), bits(qkv[0]))      # no rotation
        assert torch.equal(bits(kcs[0].cpu()), bits(ref_kc)) and torch.equal(bits(vcs[0].cpu()), bits(ref_vc))
        o, r = cap.out_host[0].float(), out.float()
        assert torch.allclose(o, r, atol=1e-3, rtol=1e-3), f"step {step}: {(o - r).abs().max()}"
