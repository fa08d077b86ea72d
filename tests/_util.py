"""Shared helpers for the test-suite (golden fixtures, reference-kernel loader)."""
import importlib.util
import json
import os

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
DT = {"float32": torch.float32, "float16": torch.float16, "bfloat16": torch.bfloat16}


def _from_raw(arr: np.ndarray, dtype: torch.dtype) -> torch.Tensor:
    if dtype in (torch.float16, torch.bfloat16):
        return torch.from_numpy(arr.view(np.int16).copy()).view(dtype)
    return torch.from_numpy(arr.copy())


def load_attention_golden():
    """List of dicts with torch CPU tensors: q, kc, vc, bt, cl, out (reference output), alibi, meta."""
    meta = json.load(open(os.path.join(GOLDEN, "attention_ref.json")))
    z = np.load(os.path.join(GOLDEN, "attention_ref.npz"))
    cases = []
    for m in meta:
        dt = DT[m["dtype"]]
        n = m["name"]
        c = dict(meta=m, q=_from_raw(z[f"{n}.q"], dt), kc=_from_raw(z[f"{n}.kc"], dt), vc=_from_raw(z[f"{n}.vc"], dt),
                 bt=torch.from_numpy(z[f"{n}.bt"].copy()), cl=torch.from_numpy(z[f"{n}.cl"].copy()),
                 out=_from_raw(z[f"{n}.out"], dt),
                 alibi=torch.from_numpy(z[f"{n}.alibi"].copy()) if m["alibi"] else None)
        cases.append(c)
    return cases


_ref_mod = None


def load_ref_cuda():
    """The UNMODIFIED reference CUDA kernels built by oracle/build_ref.sh (None if not built)."""
    global _ref_mod
    if _ref_mod is None:
        path = os.path.join(ROOT, "oracle", "_ref", "vllm_ref_C.so")
        if not os.path.exists(path):
            return None
        spec = importlib.util.spec_from_file_location("vllm_ref_C", path)
        mod = importlib.util.module_from_spec(spec)
        try:
            spec.loader.exec_module(mod)
        except ImportError:   # no CUDA driver on this machine (the module links libcuda for the reference all-reduce)
            if torch.cuda.is_available():
                raise
            return None
        _ref_mod = mod
    return _ref_mod


def bits(t: torch.Tensor) -> torch.Tensor:
    """Integer view for bit-exact comparison."""
    if t.dtype in (torch.float16, torch.bfloat16):
        return t.contiguous().view(torch.int16)
    if t.dtype == torch.float32:
        return t.contiguous().view(torch.int32)
    return t


def elt_size(dtype: torch.dtype) -> int:
    return torch.tensor([], dtype=dtype).element_size()


def make_attention_inputs(*, num_seqs, num_heads, head_size, block_size, dtype, kv_cache_dtype="auto", num_blocks,
                          context_lens, use_alibi=False, seed=0, device="cuda", distinct_blocks=False):
    """Inputs by the reference's recipe (tests/kernels/test_attention.py:137-173, vllm/utils.py:366-392):
    q, K, V ~ uniform(-D^-0.5, D^-0.5); block tables drawn with replacement from [0, NB) unless
    distinct_blocks; alibi slopes ~ randn."""
    import random
    from vllm_splitwise_b200.utils import create_kv_caches_with_random
    random.seed(seed)
    torch.random.manual_seed(seed)
    torch.cuda.manual_seed(seed)
    hq, hkv = num_heads
    scale = float(1.0 / (head_size ** 0.5))
    q = torch.empty(num_seqs, hq, head_size, dtype=dtype, device=device).uniform_(-scale, scale)
    alibi = torch.randn(hq, dtype=torch.float, device=device) if use_alibi else None
    max_ctx = max(context_lens)
    nblk = (max_ctx + block_size - 1) // block_size
    if distinct_blocks:
        assert num_blocks >= num_seqs * nblk
        bt = torch.randperm(num_blocks)[:num_seqs * nblk].view(num_seqs, nblk).to(torch.int32).to(device)
    else:
        bt = torch.tensor([[random.randint(0, num_blocks - 1) for _ in range(nblk)] for _ in range(num_seqs)],
                          dtype=torch.int32, device=device)
    cl = torch.tensor(context_lens, dtype=torch.int32, device=device)
    kcs, vcs = create_kv_caches_with_random(num_blocks, block_size, 1, hkv, head_size, kv_cache_dtype, dtype, seed,
                                            device)
    return dict(q=q, kc=kcs[0], vc=vcs[0], bt=bt, cl=cl, alibi=alibi, scale=scale, max_ctx=max_ctx,
                hq=hq, hkv=hkv, block_size=block_size, kv_cache_dtype=kv_cache_dtype)


def run_v1(ops, a, out=None):
    out = torch.empty_like(a["q"]) if out is None else out
    ops.paged_attention_v1(out, a["q"], a["kc"], a["vc"], a["hkv"], a["scale"], a["bt"], a["cl"], a["block_size"],
                           a["max_ctx"], a["alibi"], a["kv_cache_dtype"])
    return out


def run_v2(ops, a, partition_size=512):
    q = a["q"]
    s, h, d = q.shape
    p = (a["max_ctx"] + partition_size - 1) // partition_size
    out = torch.empty_like(q)
    tmp = torch.full((s, h, p, d), float("nan"), dtype=q.dtype, device=q.device)
    es = torch.full((s, h, p), float("nan"), dtype=torch.float32, device=q.device)
    ml = torch.full_like(es, float("nan"))
    ops.paged_attention_v2(out, es, ml, tmp, q, a["kc"], a["vc"], a["hkv"], a["scale"], a["bt"], a["cl"],
                           a["block_size"], a["max_ctx"], a["alibi"], a["kv_cache_dtype"])
    return out, es, ml, tmp


def dequantized(a):
    """(key_cache, value_cache) in the query dtype -- what the oracle attends over (the reference test
    dequantises fp8 caches the same way, test_attention.py:228-244)."""
    if a["kv_cache_dtype"] != "fp8_e5m2":
        return a["kc"], a["vc"]
    from oracle import paged_kv_oracle as O
    dt = a["q"].dtype
    return O.e5m2_to_tensor(a["kc"], dt), O.e5m2_to_tensor(a["vc"], dt)


def oracle_attention(a):
    from oracle import paged_kv_oracle as O
    kc, vc = dequantized(a)
    ref = torch.empty_like(a["q"])
    O.single_query_cached_kv_attention(ref, a["q"], a["hq"] // a["hkv"], kc, vc, a["bt"], a["cl"], a["scale"],
                                       a["alibi"])
    return ref


def load_rope_golden():
    """RotaryEmbedding._forward outputs of the reference (oracle/make_golden.py): list of dicts with CPU tensors
    pos [B, S], q/k inputs, oq/ok outputs ([B, S, heads*D]) and the reference's cos_sin_cache."""
    meta = json.load(open(os.path.join(GOLDEN, "rope_ref.json")))
    z = np.load(os.path.join(GOLDEN, "rope_ref.npz"))
    out = []
    for m in meta:
        dt, n = DT[m["dtype"]], m["name"]
        out.append(dict(meta=m, pos=torch.from_numpy(z[f"{n}.pos"].copy()),
                        **{k: _from_raw(z[f"{n}.{k}"], dt) for k in ("q", "k", "oq", "ok", "cache")}))
    return out
